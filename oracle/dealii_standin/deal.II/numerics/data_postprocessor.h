// Stand-in for the handful of deal.II types that dg2d/models/euler/pde.h of the reference names, so that this one header
// compiles here unmodified (deal.II itself is not installed).  TEST INFRASTRUCTURE: used only by oracle/ref_dg2d_roe.cc to
// pin the Euler Roe flux.  Only what pde.h needs to parse and what PDE::roe_flux<2> needs to run is provided; the
// arithmetic that matters is written as deal.II defines it (Tensor<1,dim> * Tensor<1,dim> = sum over d of a[d]*b[d],
// Tensor * scalar and Tensor +/- Tensor component by component).
#pragma once
#include <array>
#include <cmath>
#include <cstddef>
#include <iostream>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

namespace dealii {

template <typename T>
class Vector {
public:
    Vector() = default;
    explicit Vector(std::size_t n) : v_(n, T(0)) {}
    T &operator[](std::size_t i) { return v_[i]; }
    const T &operator[](std::size_t i) const { return v_[i]; }
    T &operator()(std::size_t i) { return v_[i]; }
    const T &operator()(std::size_t i) const { return v_[i]; }
    std::size_t size() const { return v_.size(); }

private:
    std::vector<T> v_;
};

template <int rank, int dim>
class Tensor;

template <int dim>
class Tensor<1, dim> {
public:
    Tensor() { for (int d = 0; d < dim; ++d) c_[d] = 0.0; }
    double &operator[](unsigned int d) { return c_[d]; }
    const double &operator[](unsigned int d) const { return c_[d]; }
    double norm_square() const
    {
        double s = c_[0] * c_[0];
        for (int d = 1; d < dim; ++d) s += c_[d] * c_[d];
        return s;
    }

private:
    double c_[dim];
};

template <int dim>
inline double operator*(const Tensor<1, dim> &a, const Tensor<1, dim> &b)
{
    double s = a[0] * b[0];
    for (int d = 1; d < dim; ++d) s += a[d] * b[d];
    return s;
}
template <int dim>
inline Tensor<1, dim> operator*(const Tensor<1, dim> &a, double f)
{
    Tensor<1, dim> r;
    for (int d = 0; d < dim; ++d) r[d] = a[d] * f;
    return r;
}
template <int dim>
inline Tensor<1, dim> operator*(double f, const Tensor<1, dim> &a) { return a * f; }
template <int dim>
inline Tensor<1, dim> operator+(const Tensor<1, dim> &a, const Tensor<1, dim> &b)
{
    Tensor<1, dim> r;
    for (int d = 0; d < dim; ++d) r[d] = a[d] + b[d];
    return r;
}
template <int dim>
inline Tensor<1, dim> operator-(const Tensor<1, dim> &a, const Tensor<1, dim> &b)
{
    Tensor<1, dim> r;
    for (int d = 0; d < dim; ++d) r[d] = a[d] - b[d];
    return r;
}

template <int dim>
class Point : public Tensor<1, dim> {};

template <typename T>
class FullMatrix {
public:
    FullMatrix() = default;
    FullMatrix(std::size_t m, std::size_t n) : n_(n), v_(m * n, T(0)) {}
    T &operator()(std::size_t i, std::size_t j) { return v_[i * n_ + j]; }
    const T &operator()(std::size_t i, std::size_t j) const { return v_[i * n_ + j]; }

private:
    std::size_t n_ = 0;
    std::vector<T> v_;
};

template <typename T, std::size_t N, std::size_t M>
using ndarray = std::array<std::array<T, M>, N>;

enum UpdateFlags { update_default = 0, update_values = 1 };

namespace DataPostprocessorInputs {
template <int dim>
struct Vector {
    std::vector<dealii::Vector<double>> solution_values;
};
}  // namespace DataPostprocessorInputs

template <int dim>
class DataPostprocessor {
public:
    virtual ~DataPostprocessor() = default;
    virtual void evaluate_vector_field(const DataPostprocessorInputs::Vector<dim> &, std::vector<Vector<double>> &) const = 0;
    virtual std::vector<std::string> get_names() const = 0;
    virtual UpdateFlags get_needed_update_flags() const = 0;
};

inline std::runtime_error ExcMessage(const std::string &s) { return std::runtime_error(s); }
inline std::runtime_error ExcInternalError() { return std::runtime_error("internal error"); }

}  // namespace dealii

#define Assert(cond, exc) do { if (!(cond)) throw (exc); } while (0)
#define AssertThrow(cond, exc) do { if (!(cond)) throw (exc); } while (0)
#define AssertDimension(a, b) do { if ((a) != (b)) throw std::runtime_error("dimension mismatch"); } while (0)
