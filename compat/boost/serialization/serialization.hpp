// compat/boost/serialization/serialization.hpp -- a small binary archive with the
// Boost.Serialization surface that dataArray.h:297-318, segments.h:124-131,
// sparsematrix.h:52-66 and serialization.h:47-67 of the reference rely on:
// access, make_nvp, make_array, operator& on arithmetic types, std::list/vector/map,
// class types with a member serialize(), and owning pointers to such classes.
// The byte format is this file's own (little-endian PODs, u64 container sizes); it is
// NOT Boost's archive format.  Boost is not installed on this image.
#ifndef L3D_COMPAT_BOOST_SERIALIZATION_HPP_
#define L3D_COMPAT_BOOST_SERIALIZATION_HPP_
#include <cstdint>
#include <cstddef>
#include <istream>
#include <ostream>
#include <list>
#include <map>
#include <vector>
#include <string>
#include <type_traits>
#include <utility>

namespace boost {
namespace serialization {
class access {
public:
    template<class Archive, class T>
    static void serialize(Archive& ar, T& t, const unsigned int v) { t.serialize(ar, v); }
};
template<class T> struct nvp { T* t; explicit nvp(T* p) : t(p) {} };
template<class T> inline nvp<T> make_nvp(const char*, T& t) { return nvp<T>(&t); }
template<class T> struct array_wrapper { T* p; std::size_t n; };
template<class T> inline array_wrapper<T> make_array(T* p, std::size_t n) { array_wrapper<T> a; a.p = p; a.n = n; return a; }
}  // namespace serialization

namespace archive {
namespace detail {
template<bool B> struct bool_ { static const bool value = B; };
}
template<class Derived, bool Loading>
class basic_binary_archive {
public:
    typedef detail::bool_<Loading> is_loading;
    typedef detail::bool_<!Loading> is_saving;
    Derived& self() { return *static_cast<Derived*>(this); }

    template<class T> Derived& operator&(const serialization::nvp<T>& w) {
        self().process(*const_cast<typename std::remove_const<T>::type*>(w.t)); return self();
    }
    template<class T> Derived& operator&(const serialization::array_wrapper<T>& a) {
        for (std::size_t i = 0; i < a.n; ++i) self().process(a.p[i]); return self();
    }
    template<class T> Derived& operator&(T& t) { self().process(t); return self(); }
    template<class T> Derived& operator<<(const T& t) { return (*this) & const_cast<T&>(t); }
    template<class T> Derived& operator>>(T& t) { return (*this) & t; }

    // arithmetic / enum
    template<class T>
    typename std::enable_if<std::is_arithmetic<T>::value || std::is_enum<T>::value>::type
    process(T& t) { self().raw(&t, sizeof(T)); }
    // class types with serialize()
    template<class T>
    typename std::enable_if<std::is_class<T>::value>::type
    process(T& t) { serialization::access::serialize(self(), t, 0u); }
    // owning pointers (default-constructed on load); 1 byte presence flag
    template<class T> void process(T*& p) {
        std::uint8_t present = (p != 0) ? 1 : 0;
        self().raw(&present, 1);
        if (!present) { if (Loading) p = 0; return; }
        if (Loading) p = new T();
        self().process(*p);
    }
    template<class T, class A> void process(std::vector<T, A>& v) {
        std::uint64_t n = v.size(); self().raw(&n, 8);
        if (Loading) v.resize(std::size_t(n));
        for (std::size_t i = 0; i < v.size(); ++i) self().process(v[i]);
    }
    template<class T, class A> void process(std::list<T, A>& l) {
        std::uint64_t n = l.size(); self().raw(&n, 8);
        if (Loading) { l.clear(); for (std::uint64_t i = 0; i < n; ++i) { l.push_back(T()); self().process(l.back()); } }
        else for (typename std::list<T, A>::iterator it = l.begin(); it != l.end(); ++it) self().process(*it);
    }
    template<class K, class V, class C, class A> void process(std::map<K, V, C, A>& m) {
        std::uint64_t n = m.size(); self().raw(&n, 8);
        if (Loading) {
            m.clear();
            for (std::uint64_t i = 0; i < n; ++i) { K k = K(); self().process(k); V v = V(); self().process(v); m.insert(m.end(), std::make_pair(k, v)); }
        } else for (typename std::map<K, V, C, A>::iterator it = m.begin(); it != m.end(); ++it) { K k = it->first; self().process(k); self().process(it->second); }
    }
    template<class A, class B> void process(std::pair<A, B>& p) { self().process(p.first); self().process(p.second); }
    void process(std::string& s) {
        std::uint64_t n = s.size(); self().raw(&n, 8);
        if (Loading) s.resize(std::size_t(n));
        if (n) self().raw(&s[0], std::size_t(n));
    }
};

class binary_oarchive : public basic_binary_archive<binary_oarchive, false> {
public:
    explicit binary_oarchive(std::ostream& os, unsigned int = 0) : os_(os) {}
    void raw(const void* p, std::size_t n) { os_.write(static_cast<const char*>(p), std::streamsize(n)); }
private:
    std::ostream& os_;
};
class binary_iarchive : public basic_binary_archive<binary_iarchive, true> {
public:
    explicit binary_iarchive(std::istream& is, unsigned int = 0) : is_(is) {}
    void raw(void* p, std::size_t n) { is_.read(static_cast<char*>(p), std::streamsize(n)); }
private:
    std::istream& is_;
};
}  // namespace archive
}  // namespace boost
#endif
