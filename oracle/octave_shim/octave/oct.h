// TEST INFRASTRUCTURE ONLY -- minimal stand-in for the liboctave / Octave-interpreter API that the
// reference's test driver uses (Convolution/src/convolution_test.cc:60-71,112-156,393,549-555):
//   Matrix(r,c), Matrix(), operator()(i,j), dim1(), dim2();
//   convn(Matrix, Matrix, convn_full | convn_same | convn_valid);
//   octave_value_list with operator()(int) and .matrix_value();
//   feval("fftconv" | "fftconv2", args, 1);  octave_main(argc, argv, embedded).
// Neither Octave nor its headers exist in this image, so the unchanged driver is compiled against
// this header.  Every function is a long-double nested-loop convolution (the algorithm of the
// reference's own Convolution/src/convolution_std.h:98-159), i.e. the driver then compares the
// library under test with the CPU oracle instead of with Octave.  'same' crops at floor(M/2) like
// Octave/MATLAB (not scipy's floor((M-1)/2)).
#ifndef ORACLE_OCTAVE_SHIM_H
#define ORACLE_OCTAVE_SHIM_H

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <ctime>
#include <map>
#include <sstream>
#include <string>
#include <vector>

class Matrix {
 public:
  Matrix() : r_(0), c_(0) {}
  Matrix(int r, int c) : r_(r > 0 ? r : 0), c_(c > 0 ? c : 0), d_((size_t)r_ * c_, 0.0) {}
  double& operator()(int i, int j) { return d_[(size_t)i * c_ + j]; }
  double operator()(int i, int j) const { return d_[(size_t)i * c_ + j]; }
  int dim1() const { return r_; }
  int dim2() const { return c_; }

 private:
  int r_, c_;
  std::vector<double> d_;
};

enum convn_type { convn_full, convn_same, convn_valid };

inline Matrix octave_shim_full(const Matrix& a, const Matrix& b) {
  const int ha = a.dim1(), wa = a.dim2(), hb = b.dim1(), wb = b.dim2();
  if (ha == 0 || wa == 0 || hb == 0 || wb == 0) return Matrix();
  Matrix out(ha + hb - 1, wa + wb - 1);
  for (int i = 0; i < out.dim1(); ++i)
    for (int j = 0; j < out.dim2(); ++j) {
      long double t = 0;
      const int k0 = i - hb + 1 > 0 ? i - hb + 1 : 0, k1 = i < ha - 1 ? i : ha - 1;
      const int l0 = j - wb + 1 > 0 ? j - wb + 1 : 0, l1 = j < wa - 1 ? j : wa - 1;
      for (int k = k0; k <= k1; ++k)
        for (int l = l0; l <= l1; ++l) t += (long double)a(k, l) * (long double)b(i - k, j - l);
      out(i, j) = (double)t;
    }
  return out;
}

inline Matrix convn(const Matrix& a, const Matrix& b, convn_type ct) {
  Matrix full = octave_shim_full(a, b);
  if (ct == convn_full) return full;
  int h, w, oh, ow;
  if (ct == convn_same) { h = a.dim1(); w = a.dim2(); oh = b.dim1() / 2; ow = b.dim2() / 2; }
  else { h = a.dim1() - b.dim1() + 1; w = a.dim2() - b.dim2() + 1; oh = b.dim1() - 1; ow = b.dim2() - 1; }
  if (h <= 0 || w <= 0) return Matrix();
  Matrix out(h, w);
  for (int i = 0; i < h; ++i)
    for (int j = 0; j < w; ++j) out(i, j) = full(i + oh, j + ow);
  return out;
}

class octave_value {
 public:
  octave_value() {}
  octave_value(const Matrix& m) : m_(m) {}
  octave_value& operator=(const Matrix& m) { m_ = m; return *this; }
  Matrix matrix_value() const { return m_; }

 private:
  Matrix m_;
};

class octave_value_list {
 public:
  octave_value& operator()(int i) { return v_[i]; }

 private:
  std::map<int, octave_value> v_;
};

// fftconv (1-D) and fftconv2 (2-D) both return the full linear convolution.
inline octave_value_list feval(const std::string& name, octave_value_list& args, int /*nargout*/) {
  octave_value_list out;
  if (name == "fftconv" || name == "fftconv2")
    out(0) = octave_shim_full(args(0).matrix_value(), args(1).matrix_value());
  else
    fprintf(stderr, "octave shim: unknown function %s\n", name.c_str());
  return out;
}

inline int octave_main(int, char**, int) { return 0; }

#endif
