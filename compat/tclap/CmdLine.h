// compat/tclap/CmdLine.h -- the TCLAP::CmdLine / TCLAP::ValueArg<T> surface used by
// main_vsfm.cpp:36-98 and main_bundler.cpp:36-94 ("-x value" / "--long value" flags).
// TCLAP is not installed on this image.
#ifndef L3D_COMPAT_TCLAP_CMDLINE_H_
#define L3D_COMPAT_TCLAP_CMDLINE_H_
#include <cstdlib>
#include <iostream>
#include <sstream>
#include <string>
#include <vector>
namespace TCLAP {
class Arg {
public:
    Arg(const std::string& flag, const std::string& name, const std::string& desc, bool req, const std::string& typeDesc)
        : flag_(flag), name_(name), desc_(desc), typeDesc_(typeDesc), required_(req), set_(false) {}
    virtual ~Arg() {}
    virtual bool assign(const std::string& text) = 0;
    bool isSet() const { return set_; }
    std::string flag_, name_, desc_, typeDesc_;
    bool required_, set_;
};
template<class T>
class ValueArg : public Arg {
public:
    ValueArg(const std::string& flag, const std::string& name, const std::string& desc, bool req, T value, const std::string& typeDesc)
        : Arg(flag, name, desc, req, typeDesc), value_(value) {}
    T& getValue() { return value_; }
    bool assign(const std::string& text) { return parse(text, value_); }
private:
    template<class U> static bool parse(const std::string& text, U& out) { std::istringstream is(text); is >> out; return !is.fail(); }
    static bool parse(const std::string& text, std::string& out) { out = text; return true; }
    static bool parse(const std::string& text, bool& out) {
        if (text == "1" || text == "true" || text == "TRUE") { out = true; return true; }
        if (text == "0" || text == "false" || text == "FALSE") { out = false; return true; }
        return false;
    }
    T value_;
};
class CmdLine {
public:
    CmdLine(const std::string& message, const char = ' ', const std::string& version = "none", bool = true) : message_(message), version_(version) {}
    void add(Arg& a) { args_.push_back(&a); }
    void add(Arg* a) { args_.push_back(a); }
    void parse(int argc, const char* const* argv) {
        for (int i = 1; i < argc; ++i) {
            std::string tok = argv[i];
            if (tok == "-h" || tok == "--help") { usage(argv[0]); std::exit(0); }
            Arg* hit = 0;
            for (size_t k = 0; k < args_.size(); ++k)
                if (tok == "-" + args_[k]->flag_ || tok == "--" + args_[k]->name_) hit = args_[k];
            if (!hit) { std::cerr << "PARSE ERROR: unknown argument " << tok << std::endl; usage(argv[0]); std::exit(1); }
            if (i + 1 >= argc || !hit->assign(argv[i + 1])) { std::cerr << "PARSE ERROR: bad or missing value for " << tok << std::endl; std::exit(1); }
            hit->set_ = true; ++i;
        }
        for (size_t k = 0; k < args_.size(); ++k)
            if (args_[k]->required_ && !args_[k]->set_) { std::cerr << "PARSE ERROR: required argument missing: -" << args_[k]->flag_ << " (--" << args_[k]->name_ << ")" << std::endl; usage(argv[0]); std::exit(1); }
    }
private:
    void usage(const char* prog) const {
        std::cerr << "USAGE: " << prog << " [options]   (" << message_ << ")" << std::endl;
        for (size_t k = 0; k < args_.size(); ++k)
            std::cerr << "  -" << args_[k]->flag_ << ", --" << args_[k]->name_ << " <" << args_[k]->typeDesc_ << ">" << (args_[k]->required_ ? " (required) " : " ") << args_[k]->desc_ << std::endl;
    }
    std::string message_, version_;
    std::vector<Arg*> args_;
};
}  // namespace TCLAP
#endif
