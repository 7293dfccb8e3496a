// Test-infrastructure shim (NOT product code): cv::Mat / cv::Vec3b / cv::imread as used at
// MTPC/sceneManagement.h:11,132-143 and MTPC/pathTracing.cpp:158. imread opens the raw
// "<name>.bgr" sidecar (int32 rows, int32 cols, rows*cols*3 BGR bytes) that
// oracle/make_texture_sidecar.py writes with cv2.imread, looked up in CWD then $MCPT_TEXDIR.
#pragma once
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>
namespace cv {
struct Vec3b {
    unsigned char v[3];
    unsigned char& operator[](int i) { return v[i]; }
};
struct Mat {
    int rows = 0, cols = 0;
    std::vector<unsigned char> data;
    bool empty() const { return data.empty(); }
    template <class T> T& at(int r, int c) { return *reinterpret_cast<T*>(&data[(size_t(r) * cols + c) * 3]); }
};
inline Mat imread(const std::string& name) {
    Mat m;
    std::string n = name;
    while (!n.empty() && (n.back() == '\r' || n.back() == '\n')) n.pop_back();
    FILE* f = std::fopen((n + ".bgr").c_str(), "rb");
    if (!f) {
        const char* d = std::getenv("MCPT_TEXDIR");
        if (d) f = std::fopen((std::string(d) + "/" + n + ".bgr").c_str(), "rb");
    }
    if (!f) return m;
    int hdr[2];
    if (std::fread(hdr, 4, 2, f) == 2) {
        m.rows = hdr[0]; m.cols = hdr[1];
        m.data.resize(size_t(m.rows) * m.cols * 3);
        if (std::fread(m.data.data(), 1, m.data.size(), f) != m.data.size()) { m.data.clear(); m.rows = m.cols = 0; }
    }
    std::fclose(f);
    return m;
}
}
