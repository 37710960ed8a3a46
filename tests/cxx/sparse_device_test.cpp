// GPU checks of gpr::SparseGaussianProcess<T> (gpr_b200/cxx/include/SparseGaussianProcess.h; needs a B200).
// Same data, constants and checks as the reference's deterministic sparse test where one exists:
//   S1 sin regression with inducing points = every second sample, jitter 0.1: the sparse prediction stays close to the
//      dense one                                                     (pattern of tests/SparseInferenceTest.cpp)
//   S2 with inducing points == dense samples and a small jitter the sparse posterior equals the dense one:
//      mean (Kx^T (K + s^2 I)^-1 Y) and operator()(x, x)             (identity of the Nystrom form at m = n)
//   S3 error behaviour of Initialize                                  include/SparseGaussianProcess.h:124-141, :219-221
//   S4 float instantiation
#include <cmath>
#include <iostream>
#include <memory>
#include <string>

#include "GaussianProcess.h"
#include "Kernel.h"
#include "SparseGaussianProcess.h"

using namespace gpr;

static int failures = 0;
#define EXPECT(cond, what)                                                   \
    do {                                                                     \
        if (!(cond)) {                                                       \
            ++failures;                                                      \
            std::cout << "[failed] " << what << " (" #cond ")" << std::endl; \
        } else {                                                             \
            std::cout << "[passed] " << what << std::endl;                   \
        }                                                                    \
    } while (0)

template <class T>
typename GaussianProcess<T>::VectorType V1(T v) {
    typename GaussianProcess<T>::VectorType x(1);
    x[0] = v;
    return x;
}

template <class T>
void Run(const char* tag, double tol_dense, double tol_equal) {
    typedef GaussianProcess<T> GP;
    typedef SparseGaussianProcess<T> SGP;
    const unsigned n = 120;
    const double noise = 0.1;
    auto f = [](double x) { return std::sin(x) + 0.3 * std::cos(2.5 * x); };
    typename Kernel<T>::Pointer k1(new GaussianKernel<T>((T)1.2, (T)1.0));
    typename GP::Pointer dense(new GP(k1));
    std::shared_ptr<SGP> sparse(new SGP(k1, (T)0.1)), same(new SGP(k1, (T)1e-6));
    dense->SetSigma((T)noise);
    sparse->SetSigma((T)noise);
    same->SetSigma((T)noise);
    for (unsigned i = 0; i < n; ++i) {
        const double x = -4.0 + 8.0 * i / n;
        dense->AddSample(V1<T>((T)x), V1<T>((T)f(x)));
        sparse->AddSample(V1<T>((T)x), V1<T>((T)f(x)));
        same->AddSample(V1<T>((T)x), V1<T>((T)f(x)));
        if (i % 2 == 0) sparse->AddInducingSample(V1<T>((T)x), V1<T>((T)f(x)));
        same->AddInducingSample(V1<T>((T)x), V1<T>((T)f(x)));
    }
    double e1 = 0, e2 = 0, e3 = 0;
    for (unsigned i = 0; i < 50; ++i) {
        const double x = -3.9 + 7.8 * i / 50;
        const double d = dense->Predict(V1<T>((T)x))[0];
        e1 = std::max(e1, std::fabs((double)sparse->Predict(V1<T>((T)x))[0] - d));
        e2 = std::max(e2, std::fabs((double)same->Predict(V1<T>((T)x))[0] - d));
        e3 = std::max(e3, std::fabs((double)(*same)(V1<T>((T)x), V1<T>((T)x)) - (double)(*dense)(V1<T>((T)x), V1<T>((T)x))));
    }
    std::cout << "    " << tag << ": |sparse - dense| " << e1 << ", m = n: mean " << e2 << ", variance " << e3 << std::endl;
    EXPECT(sparse->GetNumberOfInducingSamples() == n / 2, std::string(tag) + " S1 inducing samples counted");
    EXPECT(e1 <= tol_dense, std::string(tag) + " S1 sparse prediction close to the dense one");
    EXPECT(e2 <= tol_equal && e3 <= tol_equal, std::string(tag) + " S2 m = n reproduces the dense posterior");
}

void Errors() {
    typedef SparseGaussianProcess<double> SGP;
    Kernel<double>::Pointer k(new GaussianKernel<double>(1.0, 1.0));
    SGP a(k, 0.1);
    a.SetSigma(0.1);
    bool t1 = false, t2 = false, t3 = false;
    try { a.Initialize(); } catch (std::string&) { t1 = true; }
    a.AddInducingSample(V1<double>(0.0), V1<double>(0.0));
    a.AddInducingSample(V1<double>(1.0), V1<double>(0.5));
    try { a.Initialize(); } catch (std::string&) { t2 = true; }
    a.AddSample(V1<double>(0.0), V1<double>(0.0));
    try { a.Predict(V1<double>(0.3)); } catch (std::string&) { t3 = true; }  // m = 2 > n = 1
    EXPECT(t1, "S3 Initialize without inducing samples throws");
    EXPECT(t2, "S3 Initialize without dense samples throws");
    EXPECT(t3, "S3 more inducing than dense samples throws");
    a.AddSample(V1<double>(1.0), V1<double>(0.5));
    a.AddSample(V1<double>(2.0), V1<double>(0.1));
    const double y = a.Predict(V1<double>(0.5))[0];
    EXPECT(std::isfinite(y), "S3 usable after the missing samples are added");
    a.ClearInducingSamples();
    EXPECT(a.GetNumberOfInducingSamples() == 0, "S3 ClearInducingSamples");
}

int main() {
    try {
        Run<double>("double", 0.05, 1e-4);
        Run<float>("float", 0.05, 1e-3);
        Errors();
    } catch (std::string& s) {
        std::cout << "[failed] exception: " << s << std::endl;
        return 1;
    }
    std::cout << (failures ? "sparse_device_test: FAILED" : "sparse_device_test: all passed") << std::endl;
    return failures ? 1 : 0;
}
