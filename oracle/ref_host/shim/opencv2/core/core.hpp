/* TEST INFRASTRUCTURE.  Minimal stand-in for the OpenCV types the reference's host sources name (cv::Mat with
 * at<T>(), Vec3b, imread / imwrite) so that they compile unmodified; imread returns an empty matrix and imwrite
 * writes nothing -- the CPU twins that are timed touch neither. */
#ifndef MVS_SHIM_OPENCV_CORE
#define MVS_SHIM_OPENCV_CORE
#include <stdint.h>
#include <memory>
#include <string>
#include <vector>
typedef unsigned char uchar;   /* OpenCV's interface.h puts these in the global namespace */
typedef unsigned short ushort;
#define CV_8UC1 0
#define CV_8UC3 16
#define CV_32FC1 5
namespace cv {
struct Vec3b {
    unsigned char v[3];
    unsigned char &operator[](int i) { return v[i]; }
    const unsigned char &operator[](int i) const { return v[i]; }
};
class Mat {
public:
    int rows = 0, cols = 0;
    Mat() {}
    Mat(int r, int c, int type) : rows(r), cols(c), esz_(type == CV_8UC3 ? 3 : type == CV_32FC1 ? 4 : 1),
                                  buf_(std::make_shared<std::vector<unsigned char>>((size_t)r * c * esz_)) { data = buf_->data(); }
    template <class T> T &at(int i, int j) { return *reinterpret_cast<T *>(data + ((size_t)i * cols + j) * esz_); }
    bool empty() const { return !data; }
    unsigned char *data = nullptr;
private:
    int esz_ = 1;
    std::shared_ptr<std::vector<unsigned char>> buf_;
};
inline Mat imread(const std::string &, int = 1) { return Mat(); }
inline bool imwrite(const std::string &, const Mat &) { return true; }
} // namespace cv
#endif
