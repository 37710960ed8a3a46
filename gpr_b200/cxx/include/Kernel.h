// Host-side kernel classes of the B200 GP library: same names, constructors, parameter order,
// string forms and error strings as the reference's include/Kernel.h, so that code written against
// agiger/GPR (apps/GaussianProcessLearn.cpp:121, the tests) recompiles unchanged.
//
// What is different: a kernel is no longer called once per matrix element through a virtual
// (reference lib/GaussianProcess.cpp:392). Every kernel can *lower* itself into the flat post-order
// program of include/gprb200.h (Kernel::Lower), which the CUDA Gram kernels evaluate per tile.
// The scalar operator()/GetDerivative stay available on the host for single evaluations and are
// the arithmetic of reference include/Kernel.h:
//   GaussianKernel           :465-479   [sigma, scale]
//   GaussianExpKernel        :580-598   [sigma, scale] in log space
//   WhiteKernel              :695-713   [scale]
//   RationalQuadraticKernel  :794-808   [scale, sigma, alpha]
//   PeriodicKernel           :912-948   [scale, b, sigma]
//   SumKernel / ProductKernel:165-178 / :314-327   k1's parameter block first (:187-191)
#ifndef GPRB200_CXX_KERNEL_H
#define GPRB200_CXX_KERNEL_H

#include <cmath>
#include <cstdint>
#include <iomanip>
#include <limits>
#include <memory>
#include <sstream>
#include <string>
#include <vector>

#include <Eigen/Dense>

#include "MatrixIO.h"

namespace gpr {

enum KernelParameterType { Scalar, Exponential };

template <class TScalarType>
class KernelFactory;

// opcodes of the flat device program; numerically identical to enum gprb200_kernel_op
namespace kprog {
enum Op : int32_t { Gaussian = 1, GaussianExp = 2, White = 3, RationalQuadratic = 4, Periodic = 5, Sum = 6, Product = 7 };
}

template <class TScalarType>
class Kernel {
public:
    typedef Kernel Self;
    typedef std::shared_ptr<Self> Pointer;
    typedef Eigen::Matrix<TScalarType, Eigen::Dynamic, 1> VectorType;
    typedef std::vector<TScalarType> ParameterVectorType;
    typedef std::string StringParameterType;
    typedef std::vector<StringParameterType> StringParameterVectorType;

    virtual TScalarType operator()(const VectorType&, const VectorType&) const {
        throw std::string("Kernel: operator() is not implemented.");
    }
    virtual VectorType GetDerivative(const VectorType&, const VectorType&) const {
        throw std::string("Kernel: GetDerivative() is not implemented.");
    }

    virtual std::string ToString() const = 0;
    virtual unsigned GetNumberOfParameters() const = 0;
    virtual const StringParameterVectorType GetStringParameters() const { return m_StringParameters; }
    virtual ParameterVectorType GetParameters() const { return m_Parameters; }
    virtual void SetParameters(const ParameterVectorType& parameters) = 0;

    // Appends this kernel (post-order) to a device program: opcodes and the flat fp64 parameter
    // vector in GetParameters() order. Replaces per-element virtual dispatch.
    virtual void Lower(std::vector<int32_t>& ops, std::vector<double>& params) const = 0;

    std::string ParametersToString(const StringParameterVectorType& params) const {
        std::string out;
        for (const StringParameterType& p : params) out += p + ",";
        return out;
    }

    Kernel() { KernelFactory<TScalarType>::RegisterKernels(); }
    virtual ~Kernel() {}

    // equality as the reference defines it (include/Kernel.h:91-114): same string form, same string
    // parameters, numeric parameters within 10 eps
    virtual bool operator==(const Kernel<TScalarType>& b) const {
        if (ToString() != b.ToString()) return false;
        if (m_StringParameters != b.m_StringParameters) return false;
        if (m_Parameters.size() != b.m_Parameters.size()) return false;
        const TScalarType tol = 10 * std::numeric_limits<TScalarType>::epsilon();
        for (size_t i = 0; i < m_Parameters.size(); ++i)
            if (std::fabs(m_Parameters[i] - b.m_Parameters[i]) > tol) return false;
        return true;
    }
    virtual bool operator!=(const Kernel<TScalarType>& b) const { return !operator==(b); }

protected:
    ParameterVectorType m_Parameters;
    StringParameterVectorType m_StringParameters;  // what Save/Load and ToString() carry

    static std::string P2S(TScalarType p) {  // max-precision scalar -> string (digits10 + 1)
        std::ostringstream ss;
        ss << std::setprecision(std::numeric_limits<TScalarType>::digits10 + 1) << p;
        return ss.str();
    }
    static TScalarType S2P(const StringParameterType& s) {
        TScalarType p = TScalarType();
        std::istringstream ss(s);
        ss >> p;
        return p;
    }
    static TScalarType SquaredDistance(const VectorType& x, const VectorType& y) {
        const TScalarType r = (x - y).norm();  // the reference squares Eigen's norm()
        return r * r;
    }

private:
    Kernel(const Self&);
    void operator=(const Self&);
};

namespace detail {

// Shared plumbing of the five leaf kernels: parameter storage, string mirror, Load(), Lower().
// Derived supplies: Name(), FromParameters(p), Validate(p) (constructor-time guards), Refresh().
template <class T, class Derived, unsigned NP, int32_t OP>
class LeafKernel : public Kernel<T> {
public:
    typedef Kernel<T> Superclass;
    typedef std::shared_ptr<Derived> Pointer;
    typedef typename Superclass::VectorType VectorType;
    typedef typename Superclass::ParameterVectorType ParameterVectorType;
    typedef typename Superclass::StringParameterType StringParameterType;
    typedef typename Superclass::StringParameterVectorType StringParameterVectorType;

    virtual std::string ToString() const {
        return std::string(Derived::Name()) + "(" + Superclass::ParametersToString(this->m_StringParameters) + ")";
    }
    virtual unsigned GetNumberOfParameters() const { return NP; }

    virtual void SetParameters(const ParameterVectorType& parameters) {
        if (parameters.size() != NP) throw std::string(Derived::Name()) + "::SetParameters: wrong number of parameters.";
        Install(parameters, false);
    }

    static Pointer Load(const StringParameterVectorType& parameters) {
        if (parameters.size() != NP) throw std::string(Derived::Name()) + "::Load: wrong number of kernel parameters.";
        ParameterVectorType p;
        for (const StringParameterType& s : parameters) p.push_back(Superclass::S2P(s));
        return Pointer(Derived::FromParameters(p));
    }

    virtual void Lower(std::vector<int32_t>& ops, std::vector<double>& params) const {
        ops.push_back(OP);
        for (T v : this->m_Parameters) params.push_back(static_cast<double>(v));
    }

protected:
    LeafKernel() : Superclass() {}
    // stores the parameters, refreshes derived constants and the string mirror. The reference tests
    // its zero-parameter guards against the values held BEFORE assignment (include/Kernel.h:529-533,
    // :1003-1008), so they can only fire from a constructor; the same holds here.
    void Install(const ParameterVectorType& parameters, bool constructing) {
        if (constructing) static_cast<Derived*>(this)->Validate(parameters);
        this->m_Parameters = parameters;
        static_cast<Derived*>(this)->Refresh();
        this->m_StringParameters.clear();
        for (T v : parameters) this->m_StringParameters.push_back(Superclass::P2S(v));
    }
};

}  // namespace detail

// ---- GaussianKernel: k = scale^2 exp(-0.5 r^2 / sigma^2) ------------------------------------------
template <class TScalarType>
class GaussianKernel : public detail::LeafKernel<TScalarType, GaussianKernel<TScalarType>, 2, kprog::Gaussian> {
public:
    typedef detail::LeafKernel<TScalarType, GaussianKernel<TScalarType>, 2, kprog::Gaussian> Base;
    typedef Kernel<TScalarType> Superclass;
    typedef GaussianKernel Self;
    typedef std::shared_ptr<Self> Pointer;
    typedef typename Superclass::VectorType VectorType;
    typedef typename Superclass::ParameterVectorType ParameterVectorType;
    typedef typename Superclass::StringParameterType StringParameterType;

    GaussianKernel(TScalarType sigma, TScalarType scale = 1) { this->Install(ParameterVectorType{sigma, scale}, true); }
    GaussianKernel(const StringParameterType& p1, const StringParameterType& p2)
        : GaussianKernel(Superclass::S2P(p1), Superclass::S2P(p2)) {}
    virtual ~GaussianKernel() {}

    virtual TScalarType operator()(const VectorType& x, const VectorType& y) const {
        return m_Scale2 * std::exp(-0.5 * Superclass::SquaredDistance(x, y) / m_Sigma2);
    }
    virtual VectorType GetDerivative(const VectorType& x, const VectorType& y) const {
        const TScalarType r2 = Superclass::SquaredDistance(x, y);
        const TScalarType f = std::exp(-0.5 * r2 / m_Sigma2);
        VectorType D = VectorType::Zero(2);
        D[0] = m_Scale2 * r2 / m_Sigma3 * f;  // d/d sigma
        D[1] = 2 * this->m_Parameters[1] * f;  // d/d scale
        return D;
    }

    static const char* Name() { return "GaussianKernel"; }
    static Self* FromParameters(const ParameterVectorType& p) { return new Self(p[0], p[1]); }
    void Validate(const ParameterVectorType& p) const {
        if (p[0] == 0) throw std::string("GaussianKernel: sigma has to be positive");
        if (p[1] == 0) throw std::string("GaussianKernel: scale has to be positive");
    }
    void Refresh() {
        const TScalarType sigma = this->m_Parameters[0], scale = this->m_Parameters[1];
        m_Sigma2 = sigma * sigma;
        m_Sigma3 = sigma * sigma * sigma;
        m_Scale2 = scale * scale;
    }

protected:
    TScalarType m_Sigma2, m_Sigma3, m_Scale2;
};

// ---- GaussianExpKernel: Gaussian kernel with both parameters in log space ---------------------------
template <class TScalarType>
class GaussianExpKernel : public detail::LeafKernel<TScalarType, GaussianExpKernel<TScalarType>, 2, kprog::GaussianExp> {
public:
    typedef Kernel<TScalarType> Superclass;
    typedef GaussianExpKernel Self;
    typedef std::shared_ptr<Self> Pointer;
    typedef typename Superclass::VectorType VectorType;
    typedef typename Superclass::ParameterVectorType ParameterVectorType;
    typedef typename Superclass::StringParameterType StringParameterType;

    GaussianExpKernel(TScalarType sigma, TScalarType scale = 1) { this->Install(ParameterVectorType{sigma, scale}, true); }
    GaussianExpKernel(const StringParameterType& p1, const StringParameterType& p2)
        : GaussianExpKernel(Superclass::S2P(p1), Superclass::S2P(p2)) {}
    virtual ~GaussianExpKernel() {}

    virtual TScalarType operator()(const VectorType& x, const VectorType& y) const {
        const TScalarType es = std::exp(this->m_Parameters[1]), eg = std::exp(this->m_Parameters[0]);
        return es * es * std::exp(-0.5 * Superclass::SquaredDistance(x, y) / (eg * eg));
    }
    virtual VectorType GetDerivative(const VectorType& x, const VectorType& y) const {
        const TScalarType sigma = this->m_Parameters[0], scale = this->m_Parameters[1];
        const TScalarType r2 = Superclass::SquaredDistance(x, y);
        const TScalarType shrink = std::exp(-2 * sigma), grow = std::exp(2 * sigma);
        VectorType D = VectorType::Zero(2);
        D[0] = r2 * std::exp(-0.5 * shrink * ((4 * sigma - 4 * scale) * grow + r2));
        D[1] = 2 * std::exp(0.5 * shrink * (4 * grow * scale - r2));
        return D;
    }

    static const char* Name() { return "GaussianExpKernel"; }
    static Self* FromParameters(const ParameterVectorType& p) { return new Self(p[0], p[1]); }
    void Validate(const ParameterVectorType&) const {}
    void Refresh() {}
};

// ---- WhiteKernel: scale^2 where x and y coincide (||x-y|| == 0, not i == j) ---------------------------
template <class TScalarType>
class WhiteKernel : public detail::LeafKernel<TScalarType, WhiteKernel<TScalarType>, 1, kprog::White> {
public:
    typedef Kernel<TScalarType> Superclass;
    typedef WhiteKernel Self;
    typedef std::shared_ptr<Self> Pointer;
    typedef typename Superclass::VectorType VectorType;
    typedef typename Superclass::ParameterVectorType ParameterVectorType;
    typedef typename Superclass::StringParameterType StringParameterType;

    WhiteKernel(TScalarType scale) { this->Install(ParameterVectorType{scale}, true); }
    WhiteKernel(const StringParameterType& p1) : WhiteKernel(Superclass::S2P(p1)) {}
    virtual ~WhiteKernel() {}

    virtual TScalarType operator()(const VectorType& x, const VectorType& y) const {
        const TScalarType s = this->m_Parameters[0];
        return ((x - y).norm() == 0) ? s * s : TScalarType(0);
    }
    virtual VectorType GetDerivative(const VectorType& x, const VectorType& y) const {
        VectorType D = VectorType::Zero(1);
        D[0] = ((x - y).norm() == 0) ? 2 * this->m_Parameters[0] : TScalarType(0);
        return D;
    }

    static const char* Name() { return "WhiteKernel"; }
    static Self* FromParameters(const ParameterVectorType& p) { return new Self(p[0]); }
    void Validate(const ParameterVectorType&) const {}
    void Refresh() {}
};

// ---- RationalQuadraticKernel: k = scale^2 (1 + r^2 / (2 alpha sigma^2))^-alpha ---------------------------
template <class TScalarType>
class RationalQuadraticKernel
    : public detail::LeafKernel<TScalarType, RationalQuadraticKernel<TScalarType>, 3, kprog::RationalQuadratic> {
public:
    typedef Kernel<TScalarType> Superclass;
    typedef RationalQuadraticKernel Self;
    typedef std::shared_ptr<Self> Pointer;
    typedef typename Superclass::VectorType VectorType;
    typedef typename Superclass::ParameterVectorType ParameterVectorType;
    typedef typename Superclass::StringParameterType StringParameterType;

    RationalQuadraticKernel(TScalarType scale, TScalarType sigma, TScalarType alpha) {
        this->Install(ParameterVectorType{scale, sigma, alpha}, true);
    }
    RationalQuadraticKernel(const StringParameterType& p1, const StringParameterType& p2, const StringParameterType& p3)
        : RationalQuadraticKernel(Superclass::S2P(p1), Superclass::S2P(p2), Superclass::S2P(p3)) {}
    virtual ~RationalQuadraticKernel() {}

    virtual TScalarType operator()(const VectorType& x, const VectorType& y) const {
        const TScalarType alpha = this->m_Parameters[2];
        return m_Scale2 * std::pow(1 + 0.5 * Superclass::SquaredDistance(x, y) / (m_Sigma2 * alpha), -alpha);
    }
    virtual VectorType GetDerivative(const VectorType& x, const VectorType& y) const {
        const TScalarType scale = this->m_Parameters[0], alpha = this->m_Parameters[2];
        const TScalarType r2 = Superclass::SquaredDistance(x, y);
        const TScalarType base = 0.5 * r2 / (m_Sigma2 * alpha) + 1;
        const TScalarType pw = std::pow(base, -alpha);
        VectorType D = VectorType::Zero(3);
        D[0] = 2 * scale * pw;
        D[1] = m_Scale2 * r2 * std::pow(base, -alpha - 1) / m_Sigma3;
        D[2] = m_Scale2 * ((r2 / (2 * m_Sigma2 * base * alpha)) - std::log(base)) * pw;
        return D;
    }

    static const char* Name() { return "RationalQuadraticKernel"; }
    static Self* FromParameters(const ParameterVectorType& p) { return new Self(p[0], p[1], p[2]); }
    void Validate(const ParameterVectorType&) const {}
    void Refresh() {
        const TScalarType scale = this->m_Parameters[0], sigma = this->m_Parameters[1];
        m_Scale2 = scale * scale;
        m_Sigma2 = sigma * sigma;
        m_Sigma3 = sigma * sigma * sigma;
    }

protected:
    TScalarType m_Scale2, m_Sigma2, m_Sigma3;
};

// ---- PeriodicKernel: k = scale^2 exp(-0.5 sum_d sin^2(b (x_d - y_d)) / sigma^2), b = pi / period ---------
template <class TScalarType>
class PeriodicKernel : public detail::LeafKernel<TScalarType, PeriodicKernel<TScalarType>, 3, kprog::Periodic> {
public:
    typedef Kernel<TScalarType> Superclass;
    typedef PeriodicKernel Self;
    typedef std::shared_ptr<Self> Pointer;
    typedef typename Superclass::VectorType VectorType;
    typedef typename Superclass::ParameterVectorType ParameterVectorType;
    typedef typename Superclass::StringParameterType StringParameterType;

    PeriodicKernel(TScalarType scale, TScalarType b, TScalarType sigma) {
        this->Install(ParameterVectorType{scale, b, sigma}, true);
    }
    PeriodicKernel(const StringParameterType& p1, const StringParameterType& p2, const StringParameterType& p3)
        : PeriodicKernel(Superclass::S2P(p1), Superclass::S2P(p2), Superclass::S2P(p3)) {}
    virtual ~PeriodicKernel() {}

    virtual TScalarType operator()(const VectorType& x, const VectorType& y) const {
        TScalarType sin2, unused;
        Geometry(x, y, sin2, unused);
        return m_Scale2 * std::exp(-0.5 * sin2 / m_Sigma2);
    }
    virtual VectorType GetDerivative(const VectorType& x, const VectorType& y) const {
        TScalarType sin2, slope;
        Geometry(x, y, sin2, slope);
        const TScalarType e = std::exp(-0.5 * sin2 / m_Sigma2);
        VectorType D = VectorType::Zero(3);
        D[0] = 2 * this->m_Parameters[0] * e;
        D[1] = -0.5 * m_Scale2 * e * slope / m_Sigma2;
        D[2] = m_Scale2 * e * sin2 / m_Sigma3;
        return D;
    }

    static const char* Name() { return "PeriodicKernel"; }
    static Self* FromParameters(const ParameterVectorType& p) { return new Self(p[0], p[1], p[2]); }
    void Validate(const ParameterVectorType& p) const {
        if (p[0] == 0) throw std::string("PeriodicKernel: scale parameter has to be positive.");
        if (p[1] == 0) throw std::string("PeriodicKernel: period length parameter has to be positive.");
        if (p[2] == 0) throw std::string("PeriodicKernel: sigma parameter has to be positive.");
    }
    void Refresh() {
        const TScalarType scale = this->m_Parameters[0], sigma = this->m_Parameters[2];
        m_Scale2 = scale * scale;
        m_Sigma2 = sigma * sigma;
        m_Sigma3 = sigma * sigma * sigma;
    }

protected:
    // sin2 = sum_d sin^2(b dx_d), slope = sum_d 2 dx_d cos(b dx_d) sin(b dx_d); the per-dimension
    // terms are formed in double and accumulated in T as the reference does (include/Kernel.h:914-940)
    void Geometry(const VectorType& x, const VectorType& y, TScalarType& sin2, TScalarType& slope) const {
        const TScalarType b = this->m_Parameters[1];
        sin2 = 0;
        slope = 0;
        for (unsigned i = 0; i < static_cast<unsigned>(x.rows()); ++i) {
            const double dx = x[i] - y[i];
            const double s = std::sin(b * dx);
            sin2 += s * s;
            slope += 2 * dx * std::cos(b * dx) * s;
        }
    }
    TScalarType m_Scale2, m_Sigma2, m_Sigma3;
};

namespace detail {

// Sum and Product share everything but the combination rule.
template <class T, bool IS_PRODUCT>
class PairKernel : public Kernel<T> {
public:
    typedef Kernel<T> Superclass;
    typedef typename Superclass::Pointer SuperclassPointer;
    typedef typename Superclass::VectorType VectorType;
    typedef typename Superclass::ParameterVectorType ParameterVectorType;
    typedef typename Superclass::StringParameterType StringParameterType;
    typedef typename Superclass::StringParameterVectorType StringParameterVectorType;

    static const char* Name() { return IS_PRODUCT ? "ProductKernel" : "SumKernel"; }

    virtual T operator()(const VectorType& x, const VectorType& y) const {
        const T a = (*m_Kernel1)(x, y), b = (*m_Kernel2)(x, y);
        return IS_PRODUCT ? a * b : a + b;
    }
    // [dk1 ; dk2] for a sum, [dk1 * k2 ; dk2 * k1] for a product
    virtual VectorType GetDerivative(const VectorType& x, const VectorType& y) const {
        VectorType D1 = m_Kernel1->GetDerivative(x, y);
        VectorType D2 = m_Kernel2->GetDerivative(x, y);
        if (IS_PRODUCT) {
            const T a = (*m_Kernel1)(x, y), b = (*m_Kernel2)(x, y);
            D1 = D1 * b;
            D2 = D2 * a;
        }
        VectorType D = VectorType::Zero(D1.rows() + D2.rows());
        D.topRows(D1.rows()) = D1;
        D.bottomRows(D2.rows()) = D2;
        return D;
    }

    virtual std::string ToString() const {
        return std::string(Name()) + "(" + m_Kernel1->ToString() + "," + m_Kernel2->ToString() + ")";
    }
    virtual unsigned GetNumberOfParameters() const {
        return m_Kernel1->GetNumberOfParameters() + m_Kernel2->GetNumberOfParameters();
    }
    virtual void SetParameters(const ParameterVectorType& parameters) {
        if (parameters.size() != GetNumberOfParameters())
            throw std::string(Name()) + "::SetParameters: wrong number of parameters.";
        Distribute(parameters);
    }
    virtual void Lower(std::vector<int32_t>& ops, std::vector<double>& params) const {
        m_Kernel1->Lower(ops, params);
        m_Kernel2->Lower(ops, params);
        ops.push_back(IS_PRODUCT ? kprog::Product : kprog::Sum);
    }

    const SuperclassPointer GetKernel1() { return m_Kernel1; }
    const SuperclassPointer GetKernel2() { return m_Kernel2; }

protected:
    PairKernel(SuperclassPointer k1, SuperclassPointer k2) : Superclass(), m_Kernel1(k1), m_Kernel2(k2) {
        ParameterVectorType all = k1->GetParameters();
        const ParameterVectorType second = k2->GetParameters();
        all.insert(all.end(), second.begin(), second.end());
        Distribute(all);
    }

    // pushes the two parameter blocks down to the children and rebuilds the string mirror:
    // [k1 string, k2 string, #string-params of k1, of k2, k1's string params..., k2's ...]
    void Distribute(const ParameterVectorType& parameters) {
        const unsigned n1 = m_Kernel1->GetNumberOfParameters();
        m_Kernel1->SetParameters(ParameterVectorType(parameters.begin(), parameters.begin() + n1));
        m_Kernel2->SetParameters(ParameterVectorType(parameters.begin() + n1, parameters.end()));
        this->m_Parameters = parameters;
        const StringParameterVectorType s1 = m_Kernel1->GetStringParameters(), s2 = m_Kernel2->GetStringParameters();
        StringParameterVectorType& sp = this->m_StringParameters;
        sp.clear();
        sp.push_back(m_Kernel1->ToString());
        sp.push_back(m_Kernel2->ToString());
        sp.push_back(Superclass::P2S(static_cast<T>(s1.size())));
        sp.push_back(Superclass::P2S(static_cast<T>(s2.size())));
        sp.insert(sp.end(), s1.begin(), s1.end());
        sp.insert(sp.end(), s2.begin(), s2.end());
    }

    // children are rebuilt from their ToString() forms (the first two string parameters)
    static void LoadChildren(const StringParameterVectorType& parameters, SuperclassPointer& k1, SuperclassPointer& k2) {
        if (parameters.size() < 4) throw std::string(Name()) + "::Load: wrong number of kernel parameters.";
        std::string s1 = parameters[0], s2 = parameters[1];
        k1 = KernelFactory<T>::GetKernel(s1);
        k2 = KernelFactory<T>::GetKernel(s2);
    }

    SuperclassPointer m_Kernel1;
    SuperclassPointer m_Kernel2;
};

}  // namespace detail

// ---- SumKernel: k(x,y) = k1(x,y) + k2(x,y) ---------------------------------------------------------
template <class TScalarType>
class SumKernel : public detail::PairKernel<TScalarType, false> {
public:
    typedef Kernel<TScalarType> Superclass;
    typedef typename Superclass::Pointer SuperclassPointer;
    typedef SumKernel Self;
    typedef std::shared_ptr<Self> Pointer;
    typedef typename Superclass::StringParameterVectorType StringParameterVectorType;

    SumKernel(SuperclassPointer const k1, SuperclassPointer const k2) : detail::PairKernel<TScalarType, false>(k1, k2) {}
    virtual ~SumKernel() {}
    static Pointer Load(const StringParameterVectorType& parameters) {
        SuperclassPointer k1, k2;
        Self::LoadChildren(parameters, k1, k2);
        return Pointer(new Self(k1, k2));
    }
};

// ---- ProductKernel: k(x,y) = k1(x,y) * k2(x,y) -------------------------------------------------------
template <class TScalarType>
class ProductKernel : public detail::PairKernel<TScalarType, true> {
public:
    typedef Kernel<TScalarType> Superclass;
    typedef typename Superclass::Pointer SuperclassPointer;
    typedef ProductKernel Self;
    typedef std::shared_ptr<Self> Pointer;
    typedef typename Superclass::StringParameterVectorType StringParameterVectorType;

    ProductKernel(SuperclassPointer const k1, SuperclassPointer const k2) : detail::PairKernel<TScalarType, true>(k1, k2) {}
    virtual ~ProductKernel() {}
    static Pointer Load(const StringParameterVectorType& parameters) {
        SuperclassPointer k1, k2;
        Self::LoadChildren(parameters, k1, k2);
        return Pointer(new Self(k1, k2));
    }
};

}  // namespace gpr

#include "KernelFactory.h"

#endif
