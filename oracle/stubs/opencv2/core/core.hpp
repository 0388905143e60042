// Stand-in for <opencv2/core/core.hpp>, for building the reference's DBoW2 sources in place (oracle/Makefile,
// _ref/libvisnav_ref_bow.so).  TEST INFRASTRUCTURE ONLY.  DBoW2 uses cv::Mat as a reference-counted byte row
// (create / zeros / ptr / rows / cols / release / clone) and names cv::FileStorage / cv::FileNode in its YAML
// save / load, which the oracle never calls: those two classes only have to let the templates compile.
#pragma once
#include <math.h>

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <sstream>
#include <memory>
#include <string>
#include <vector>

#define CV_8U 0
#define CV_32F 5

namespace cv {

class Mat {
 public:
  int rows = 0, cols = 0;
  Mat() {}
  void create(int r, int c, int type) {
    esz_ = type == CV_32F ? 4 : 1;
    rows = r;
    cols = c;
    buf_ = std::make_shared<std::vector<unsigned char>>((size_t)r * c * esz_);
  }
  static Mat zeros(int r, int c, int type) {
    Mat m;
    m.create(r, c, type);
    return m;  // vector storage is value-initialised
  }
  Mat clone() const {
    Mat m;
    m.rows = rows; m.cols = cols; m.esz_ = esz_;
    if (buf_) m.buf_ = std::make_shared<std::vector<unsigned char>>(*buf_);
    return m;
  }
  void release() { buf_.reset(); rows = cols = 0; }
  bool empty() const { return !buf_ || buf_->empty(); }
  template <class T> T* ptr(int r = 0) { return reinterpret_cast<T*>(buf_->data() + (size_t)r * cols * esz_); }
  template <class T> const T* ptr(int r = 0) const { return reinterpret_cast<const T*>(buf_->data() + (size_t)r * cols * esz_); }

 private:
  std::shared_ptr<std::vector<unsigned char>> buf_;
  int esz_ = 1;
};

class FileNode {
 public:
  enum { SEQ = 4 };
  int type() const { return 0; }
  size_t size() const { return 0; }
  FileNode operator[](const char*) const { return FileNode(); }
  FileNode operator[](const std::string&) const { return FileNode(); }
  FileNode operator[](int) const { return FileNode(); }
  operator int() const { return 0; }
  operator double() const { return 0.0; }
  operator std::string() const { return std::string(); }
};

class FileStorage {
 public:
  enum { READ = 0, WRITE = 1 };
  FileStorage(const char*, int) {}
  bool isOpened() const { return false; }
  FileNode operator[](const std::string&) const { return FileNode(); }
  FileNode operator[](const char*) const { return FileNode(); }
};
template <class T> inline FileStorage& operator<<(FileStorage& f, const T&) { return f; }

}  // namespace cv
