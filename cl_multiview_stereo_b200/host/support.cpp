// support.cpp -- helpers of the drop-in host API: the cl::Program stand-in, the vec helpers
// and message printers declared in include/header.h (reference header.h:40-53,
// file_handler.cpp:88-181) and the image I/O of include/file_handler.h (reference
// file_handler.cpp:6-85) on binary PPM instead of OpenCV.
#include "file_handler.h"

#include <stdexcept>

namespace cl {
Program::Program(int device)
{
    auto h = std::make_shared<Holder>();
    if (mvs_ctx_create(device, &h->ctx))
        throw std::runtime_error(std::string("cl::Program (CUDA context): ") + mvs_last_error());
    h_ = h;
}
} // namespace cl

void printText(std::string txt) { std::cout << txt << std::endl; }

void printError(cl_int err_no, std::string message)
{
    std::cout << "Error No " << err_no << ": " << message << std::endl;
}

// The reference's switch falls through every case (file_handler.cpp:97-113); one message per
// failure here, and the caller stops instead of continuing.
void errorHandler(CL_Error_TYPE type, cl_int err, std::string msg)
{
    if (err == CL_SUCCESS) return;
    static const char *what[] = {"Buffer Allocation Error.", "Kernel Definition Error.", "Kernel Execution Error.",
                                 "Read From Device to Host Error."};
    printError(err, std::string(what[(int)type]) + msg);
}

vec2f makeVec2f(float x, float y) { vec2f t; t.x = x; t.y = y; return t; }
vec3f makeVec3f(float x, float y, float z) { vec3f t; t.x = x; t.y = y; t.z = z; t.w = 0.0f; return t; }
vec2i makeVec2i(int x, int y) { vec2i t; t.x = x; t.y = y; return t; }
vec3i makeVec3i(int x, int y, int z) { vec3i t; t.x = x; t.y = y; t.z = z; t.w = 0; return t; }
vec8f makeVec8f(float s0, float s1, float s2, float s3, float s4, float s5, float s6, float s7)
{
    vec8f t;
    const float v[8] = {s0, s1, s2, s3, s4, s5, s6, s7};
    for (int k = 0; k < 8; ++k) t.s[k] = v[k];
    return t;
}

float euDistance3D(vec3f d1, vec3f d2)
{
    float dx = d1.x - d2.x, dy = d1.y - d2.y, dz = d1.z - d2.z;
    return sqrtf(dx * dx + dy * dy + dz * dz);
}

// the mathematically correct cross product (the reference's host helper has a wrong first
// component, file_handler.cpp:167; its device helper clcode.cl:1680-1682 is correct)
vec3f crossVec3f(vec3f a, vec3f b)
{
    return makeVec3f(a.y * b.z - a.z * b.y, b.x * a.z - a.x * b.z, a.x * b.y - a.y * b.x);
}

vec3f normalizeVec3f(vec3f a)
{
    float n = sqrtf(a.x * a.x + a.y * a.y + a.z * a.z);
    return makeVec3f(a.x / n, a.y / n, a.z / n);
}

// ---- image I/O -----------------------------------------------------------------------------
void loadImageIn(Mat &in_cam, vec3u *in_img, int height, int width)
{
    for (int i = 0; i < height; i++)
        for (int j = 0; j < width; j++) {
            const unsigned char *p = in_cam.ptr(i, j);   // B G R
            vec3u &o = in_img[(size_t)i * width + j];
            o.s0 = p[2]; o.s1 = p[1]; o.s2 = p[0]; o.s3 = 0;
        }
}

void loadImageOut(Mat &out_cam, cl_uchar3 *out_img, int height, int width)
{
    if (out_cam.rows != height || out_cam.cols != width) {
        out_cam.rows = height; out_cam.cols = width;
        out_cam.data.assign((size_t)height * width * 3, 0);
    }
    for (int i = 0; i < height; i++)
        for (int j = 0; j < width; j++) {
            unsigned char *p = out_cam.ptr(i, j);
            const cl_uchar3 &o = out_img[(size_t)i * width + j];
            p[2] = o.s0; p[1] = o.s1; p[0] = o.s2;
        }
}

static bool ppm_token(std::istream &in, int &v)
{
    for (;;) {
        int c = in.peek();
        if (c == '#') { std::string skip; std::getline(in, skip); }
        else if (c == ' ' || c == '\n' || c == '\r' || c == '\t') in.get();
        else break;
    }
    return (bool)(in >> v);
}

bool read_ppm(const std::string &path, Mat &out)
{
    std::ifstream in(path, std::ios::binary);
    if (!in) return false;
    char magic[2];
    in.read(magic, 2);
    if (!in || magic[0] != 'P' || magic[1] != '6') return false;
    int w = 0, h = 0, mx = 0;
    if (!ppm_token(in, w) || !ppm_token(in, h) || !ppm_token(in, mx) || mx != 255 || w <= 0 || h <= 0) return false;
    in.get();   // single whitespace after maxval
    std::vector<unsigned char> rgb((size_t)w * h * 3);
    in.read((char *)rgb.data(), (std::streamsize)rgb.size());
    if ((size_t)in.gcount() != rgb.size()) return false;
    out.rows = h; out.cols = w;
    out.data.resize(rgb.size());
    for (size_t i = 0; i < (size_t)w * h; ++i) {   // PPM is RGB, the raster is BGR like cv::imread
        out.data[3 * i + 0] = rgb[3 * i + 2];
        out.data[3 * i + 1] = rgb[3 * i + 1];
        out.data[3 * i + 2] = rgb[3 * i + 0];
    }
    return true;
}

bool write_ppm(const std::string &path, const Mat &img)
{
    std::ofstream o(path, std::ios::binary);
    if (!o) return false;
    o << "P6\n" << img.cols << " " << img.rows << "\n255\n";
    std::vector<unsigned char> rgb(img.data.size());
    for (size_t i = 0; i < img.data.size() / 3; ++i) {
        rgb[3 * i + 0] = img.data[3 * i + 2];
        rgb[3 * i + 1] = img.data[3 * i + 1];
        rgb[3 * i + 2] = img.data[3 * i + 0];
    }
    o.write((const char *)rgb.data(), (std::streamsize)rgb.size());
    return (bool)o;
}

bool write_pgm(const std::string &path, const unsigned char *gray, int width, int height)
{
    std::ofstream o(path, std::ios::binary);
    if (!o) return false;
    o << "P5\n" << width << " " << height << "\n255\n";
    o.write((const char *)gray, (std::streamsize)width * height);
    return (bool)o;
}

bool read_image_array(vector<Mat> &img_array, std::string img_file_addr, int num_view)
{
    std::ifstream info_file(img_file_addr);
    if (!info_file.is_open()) return true;
    // paths are relative to the list file's directory unless absolute
    std::string dir;
    size_t slash = img_file_addr.find_last_of('/');
    if (slash != std::string::npos) dir = img_file_addr.substr(0, slash + 1);
    vector<string> img_files;
    string line;
    while (getline(info_file, line)) {
        while (!line.empty() && (line.back() == '\r' || line.back() == ' ')) line.pop_back();
        if (!line.empty()) img_files.push_back(line);
    }
    if ((int)img_files.size() < num_view) return true;
    for (int i = 0; i < num_view; i++) {
        Mat m;
        std::string p = img_files[i];
        if (!read_ppm(p, m) && !(p[0] != '/' && read_ppm(dir + p, m))) {
            std::cout << "cannot read image (binary PPM expected): " << p << std::endl;
            return true;
        }
        img_array.push_back(std::move(m));
    }
    return false;
}

static std::string g_out_dir = ".";

void imgSave(Mat img, int view_no) { write_ppm(g_out_dir + "/res" + std::to_string(view_no) + ".ppm", img); }

void show_img(vec3u *res_img, int height, int width, int view_num)
{
    Mat m;
    loadImageOut(m, res_img, height, width);
    imgSave(m, view_num);
}
