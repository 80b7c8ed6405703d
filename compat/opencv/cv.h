// compat/opencv/cv.h -- shell of the OpenCV 2.x C++ names that the Line3D public API
// (line3D.h:70-80 takes a cv::Mat) and its sources/mains mention.  OpenCV's C++ headers
// are not installed on this image.  cv::Mat here is a plain owning 2-D byte buffer
// (rows, cols, type); there is NO image codec, NO resampling and NO line-segment
// detector behind these names: imread() only understands binary PGM/PPM (P5/P6) and a
// tiny "L3DSIZE <cols> <rows>" text stub that yields an image of that size, and
// LineSegmentDetector::detect() finds nothing.  Segments therefore enter through the
// segment cache (line3D.cc:160-168) -- 2-D line detection is upstream of, and outside,
// the matching-and-scoring path this repository implements.
#ifndef L3D_COMPAT_OPENCV_CV_H_
#define L3D_COMPAT_OPENCV_CV_H_
#include <cstddef>
#include <cstring>
#include <fstream>
#include <memory>
#include <sstream>
#include <string>
#include <vector>

#define CV_CN_SHIFT 3
#define CV_MAKETYPE(depth, cn) ((depth) + (((cn) - 1) << CV_CN_SHIFT))
#define CV_8U 0
#define CV_32F 5
#define CV_64F 6
#define CV_8UC1 CV_MAKETYPE(CV_8U, 1)
#define CV_8UC3 CV_MAKETYPE(CV_8U, 3)
#define CV_32FC1 CV_MAKETYPE(CV_32F, 1)
#define CV_64FC1 CV_MAKETYPE(CV_64F, 1)
#define CV_RGB2GRAY 7
#define CV_BGR2GRAY 6
#define CV_EXPORTS
#define CV_WRAP

namespace cv {

struct Size { int width, height; Size() : width(0), height(0) {} Size(int w, int h) : width(w), height(h) {} };
struct Point { int x, y; Point() : x(0), y(0) {} Point(int x_, int y_) : x(x_), y(y_) {} Point(float x_, float y_) : x(int(x_)), y(int(y_)) {} Point(double x_, double y_) : x(int(x_)), y(int(y_)) {} };
struct Scalar { double val[4]; Scalar(double a = 0, double b = 0, double c = 0, double d = 0) { val[0] = a; val[1] = b; val[2] = c; val[3] = d; } double operator[](int i) const { return val[i]; } };
template<class T, int N> struct Vec { T val[N]; Vec() { for (int i = 0; i < N; ++i) val[i] = T(0); } T& operator[](int i) { return val[i]; } const T& operator[](int i) const { return val[i]; } };
typedef Vec<float, 4> Vec4f;
typedef Vec<int, 4> Vec4i;

enum { INTER_NEAREST = 0, INTER_LINEAR = 1 };
enum { BORDER_CONSTANT = 0 };

class Mat {
public:
    int rows, cols;
    Mat() : rows(0), cols(0), type_(CV_8UC1) {}
    Mat(int r, int c, int type) : rows(r), cols(c), type_(type), buf_(new std::vector<unsigned char>(size_t(r) * c * elemSize_(type), 0)) {}
    Mat(int r, int c, int type, const Scalar& s) : rows(r), cols(c), type_(type), buf_(new std::vector<unsigned char>(size_t(r) * c * elemSize_(type), 0)) { fill(s); }
    Mat(Size sz, int type) : rows(sz.height), cols(sz.width), type_(type), buf_(new std::vector<unsigned char>(size_t(sz.height) * sz.width * elemSize_(type), 0)) {}
    static Mat zeros(int r, int c, int type) { return Mat(r, c, type); }
    int type() const { return type_; }
    int depth() const { return type_ & 7; }
    int channels() const { return (type_ >> CV_CN_SHIFT) + 1; }
    bool empty() const { return rows == 0 || cols == 0; }
    size_t elemSize() const { return elemSize_(type_); }
    Mat clone() const { Mat m(*this); if (buf_) m.buf_.reset(new std::vector<unsigned char>(*buf_)); return m; }
    unsigned char* data() { return buf_ ? &(*buf_)[0] : 0; }
    const unsigned char* data() const { return buf_ ? &(*buf_)[0] : 0; }
    template<class T> T& at(int r, int c) { return *reinterpret_cast<T*>(data() + (size_t(r) * cols + c) * elemSize()); }
    template<class T> T& at(int i) { return *reinterpret_cast<T*>(data() + size_t(i) * elemSize()); }
    template<class T> const T& at(int r, int c) const { return *reinterpret_cast<const T*>(data() + (size_t(r) * cols + c) * elemSize()); }
private:
    static size_t elemSize_(int type) { static const size_t d[8] = {1, 1, 2, 2, 4, 4, 8, 2}; return d[type & 7] * size_t((type >> CV_CN_SHIFT) + 1); }
    void fill(const Scalar& s) {
        const int cn = channels();
        for (size_t i = 0; i < size_t(rows) * cols; ++i)
            for (int k = 0; k < cn; ++k) {
                unsigned char* p = data() + (i * cn + k) * (elemSize() / cn);
                switch (depth()) { case CV_64F: *reinterpret_cast<double*>(p) = s[k]; break; case CV_32F: *reinterpret_cast<float*>(p) = float(s[k]); break; default: *p = (unsigned char)(s[k]); }
            }
    }
    int type_;
    std::shared_ptr<std::vector<unsigned char> > buf_;
};

template<class T> struct DataDepth_;
template<> struct DataDepth_<double> { enum { value = CV_64F }; };
template<> struct DataDepth_<float> { enum { value = CV_32F }; };
template<> struct DataDepth_<unsigned char> { enum { value = CV_8U }; };
template<class T> class Mat_ : public Mat {
public:
    Mat_() {}
    Mat_(int r, int c) : Mat(r, c, CV_MAKETYPE(DataDepth_<T>::value, 1)) {}
    static Mat_ zeros(int r, int c) { return Mat_(r, c); }
    static Mat_ eye(int r, int c) { Mat_ m(r, c); for (int i = 0; i < (r < c ? r : c); ++i) m.template at<T>(i, i) = T(1); return m; }
};

template<class T> class Ptr : public std::shared_ptr<T> {
public:
    Ptr() {}
    Ptr(T* p) : std::shared_ptr<T>(p) {}
};

// No detector behind this name on this image: detect() reports no segments.
enum { LSD_REFINE_NONE = 0, LSD_REFINE_STD = 1, LSD_REFINE_ADV = 2 };
class LineSegmentDetector {
public:
    virtual ~LineSegmentDetector() {}
    virtual void detect(const Mat&, std::vector<Vec4f>& lines, std::vector<double>& width,
                        std::vector<double>& prec, std::vector<double>& nfa) { lines.clear(); width.clear(); prec.clear(); nfa.clear(); }
};
inline Ptr<LineSegmentDetector> createLineSegmentDetectorPtr(int = LSD_REFINE_STD, double = 0.8, double = 0.6, double = 2.0,
                                                             double = 22.5, double = 0, double = 0.7, int = 1024) {
    return Ptr<LineSegmentDetector>(new LineSegmentDetector());
}

// geometry-only stand-ins (pixels are not resampled)
inline void resize(const Mat& src, Mat& dst, Size sz, double = 0, double = 0, int = INTER_LINEAR) { dst = Mat(sz.height, sz.width, src.type()); }
inline void cvtColor(const Mat& src, Mat& dst, int) { dst = Mat(src.rows, src.cols, CV_8UC1); }
inline void line(Mat&, Point, Point, const Scalar&, int = 1) {}
inline void initUndistortRectifyMap(const Mat&, const Mat&, const Mat&, const Mat&, Size sz, int, Mat& map1, Mat& map2) {
    map1 = Mat(sz.height, sz.width, CV_32FC1); map2 = Mat(sz.height, sz.width, CV_32FC1);
}
inline void remap(const Mat& src, Mat& dst, const Mat&, const Mat&, int, int = BORDER_CONSTANT, const Scalar& = Scalar()) { Mat s = src.clone(); dst = s; }

// binary PGM/PPM, or a one-line text stub "L3DSIZE <cols> <rows>"
inline Mat imread(const std::string& filename, int = 1) {
    std::ifstream f(filename.c_str(), std::ios::binary);
    if (!f) return Mat();
    std::string magic; f >> magic;
    if (magic == "L3DSIZE") { int c = 0, r = 0; f >> c >> r; return (c > 0 && r > 0) ? Mat(r, c, CV_8UC3) : Mat(); }
    if (magic != "P5" && magic != "P6") return Mat();
    int vals[3], got = 0;
    while (got < 3 && f) {
        f >> std::ws;
        if (f.peek() == '#') { std::string skip; std::getline(f, skip); continue; }
        f >> vals[got++];
    }
    if (got < 3 || vals[0] <= 0 || vals[1] <= 0) return Mat();
    f.get();
    Mat m(vals[1], vals[0], magic == "P5" ? CV_8UC1 : CV_8UC3);
    f.read(reinterpret_cast<char*>(m.data()), std::streamsize(size_t(m.rows) * m.cols * m.elemSize()));
    return m;
}

}  // namespace cv
#endif
