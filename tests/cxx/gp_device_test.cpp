// GPU checks of the C++ host API over libgprb200.so (needs a B200). Same constants and thresholds as the
// reference's deterministic tests where one exists:
//   T1 sin regression, Gaussian(2.889), 10 samples, sum|err| <= 8e-4   tests/GaussianProcessTest.cpp:35-76
//      under every InversionMethod value (<= 6e-4 there)                 tests/InversionMethodsTest.cpp:36-145
//   T2 2-in/2-out sin/cos, Gaussian(3.24), sum|err| <= 5e-3              tests/GaussianProcessTest.cpp:78-121
//   T3 PredictDerivative of sin ~ cos, Gaussian(1), sum|err| <= 0.6      tests/GaussianProcessTest.cpp:238-279
//   T4 GetCredibleInterval(x) == 2 sqrt(gp(x,x)) exactly                 tests/PosteriorProcessTest.cpp:51-95
//   T5 Save / Load / == incl. efficient storage, float and double        tests/IOTest.cpp:92-227
//   T6 log-likelihood gradient vs central differences of the value       tests/GaussianLikelihoodTest.cpp:147-234
//   T7 PredictBatch == Predict / GetCredibleInterval row by row          (additive API)
#include <cmath>
#include <cstdio>
#include <iostream>
#include <memory>
#include <string>

#include "GaussianProcess.h"
#include "Kernel.h"
#include "Likelihood.h"

using namespace gpr;

static int failures = 0;
#define EXPECT(cond, what)                                                                   \
    do {                                                                                     \
        if (!(cond)) {                                                                       \
            ++failures;                                                                      \
            std::cout << "[failed] " << what << " (" #cond ")" << std::endl;                 \
        } else {                                                                             \
            std::cout << "[passed] " << what << std::endl;                                   \
        }                                                                                    \
    } while (0)

template <class T>
typename GaussianProcess<T>::VectorType Vec1(T v) {
    typename GaussianProcess<T>::VectorType x(1);
    x[0] = v;
    return x;
}

template <class T>
std::shared_ptr<GaussianProcess<T> > SinModel(T sigma, unsigned n, double interval) {
    typedef GaussianProcess<T> GP;
    typename GP::Pointer gp(new GP(typename Kernel<T>::Pointer(new GaussianKernel<T>(sigma))));
    gp->SetSigma(0);
    for (unsigned i = 0; i < n; ++i) {
        const double x = i * interval / n;
        gp->AddSample(Vec1<T>((T)x), Vec1<T>((T)std::sin(x)));
    }
    return gp;
}

int main() {
    typedef GaussianProcess<double> GP;
    typedef GP::VectorType Vec;
    typedef GP::MatrixType Mat;
    try {
        // ---- T1 -----------------------------------------------------------------------------------------
        const GP::InversionMethod methods[] = {GP::FullPivotLU, GP::JacobiSVD, GP::BDCSVD, GP::SelfAdjointEigenSolver};
        for (GP::InversionMethod m : methods) {
            GP::Pointer gp = SinModel<double>(2.889, 10, 2 * M_PI);
            gp->SetInversionMethod(m);
            gp->Initialize();
            double err = 0;
            for (unsigned i = 0; i < 50; ++i) {
                const double x = i * 2 * M_PI / 50;
                err += std::fabs(gp->Predict(Vec1<double>(x))[0] - std::sin(x));
            }
            std::cout << "T1 method " << (int)m << " sum|err| = " << err << std::endl;
            EXPECT(err <= 6e-4, "T1 sin regression below the reference threshold");
        }
        // ---- T2 -----------------------------------------------------------------------------------------
        {
            GP::Pointer gp(new GP(Kernel<double>::Pointer(new GaussianKernel<double>(3.24))));
            gp->SetSigma(0);
            const unsigned n = 10;
            for (unsigned i = 0; i < n; ++i) {
                Vec x(2), y(2);
                x[0] = x[1] = i * 2 * M_PI / n;
                y[0] = std::sin(x[0]);
                y[1] = std::cos(x[0]);
                gp->AddSample(x, y);
            }
            double err = 0;
            for (unsigned i = 0; i < 50; ++i) {
                Vec x(2);
                x[0] = x[1] = i * 2 * M_PI / 50;
                const Vec y = gp->Predict(x);
                err += std::fabs(y[0] - std::sin(x[0])) + std::fabs(y[1] - std::cos(x[0]));
            }
            std::cout << "T2 sum|err| = " << err << std::endl;
            EXPECT(err <= 5e-3, "T2 two-output regression below the reference threshold");
        }
        // ---- T3 -----------------------------------------------------------------------------------------
        {
            GP::Pointer gp = SinModel<double>(1.0, 20, 4 * M_PI);
            double err = 0;
            for (unsigned i = 0; i < 50; ++i) {
                const double x = i * 4 * M_PI / 50;
                Mat D;
                const Vec y = gp->PredictDerivative(Vec1<double>(x), D);
                EXPECT(D.rows() == 1 && D.cols() == 1, "T3 derivative shape d x M");
                err += std::fabs(D(0, 0) - std::cos(x));
                if (i == 7) EXPECT(std::fabs(y[0] - gp->Predict(Vec1<double>(x))[0]) < 1e-12, "T3 mean equals Predict");
            }
            std::cout << "T3 sum|err| = " << err << std::endl;
            EXPECT(err <= 0.6, "T3 derivative of sin close to cos");
        }
        // ---- T4 -----------------------------------------------------------------------------------------
        {
            GP::Pointer gp = SinModel<double>(1.3, 12, 2 * M_PI);
            gp->SetSigma(0.1);
            bool exact = true;
            for (unsigned i = 0; i < 9; ++i) {
                const Vec x = Vec1<double>(0.37 + 0.61 * i);
                const double c = (*gp)(x, x);
                exact = exact && (gp->GetCredibleInterval(x) == 2 * std::sqrt(std::max(0.0, c)));
                exact = exact && c >= 0 && c <= 1.0 + 1e-12;  // posterior variance never exceeds the prior k(x,x) = 1
            }
            EXPECT(exact, "T4 credible interval identity");
            const Vec a = Vec1<double>(0.4), b = Vec1<double>(2.9);
            EXPECT(std::fabs((*gp)(a, b) - (*gp)(b, a)) < 1e-13, "T4 posterior covariance is symmetric");
        }
        // ---- T5 -----------------------------------------------------------------------------------------
        {
            GP::Pointer gp = SinModel<double>(1.1, 15, 2 * M_PI);
            gp->SetSigma(0.05);
            gp->Initialize();
            gp->Save("/tmp/gprb200_cxx_io");
            GP::Pointer back(new GP(Kernel<double>::Pointer(new WhiteKernel<double>(1))));
            back->Load("/tmp/gprb200_cxx_io");
            EXPECT(*gp == *back, "T5 Save/Load round trip compares equal");
            const Vec x = Vec1<double>(1.234);
            EXPECT(std::fabs(gp->Predict(x)[0] - back->Predict(x)[0]) < 1e-13, "T5 loaded model predicts identically");
            EXPECT(std::fabs((*gp)(x, x) - (*back)(x, x)) < 1e-9, "T5 loaded core matrix gives the same covariance");
            // efficient storage: no core matrix in the file, rebuilt lazily by operator()
            GP::Pointer eff = SinModel<double>(1.1, 15, 2 * M_PI);
            eff->SetSigma(0.05);
            eff->SetEfficientStorage(true);
            eff->Initialize();
            eff->Save("/tmp/gprb200_cxx_io_eff");
            FILE* f = std::fopen("/tmp/gprb200_cxx_io_eff-CoreMatrix.txt", "rb");
            std::fseek(f, 0, SEEK_END);
            EXPECT(std::ftell(f) == 4, "T5 efficient storage writes an empty '0 0' core matrix");
            std::fclose(f);
            GP::Pointer eff_back(new GP(Kernel<double>::Pointer(new WhiteKernel<double>(1))));
            eff_back->Load("/tmp/gprb200_cxx_io_eff");
            EXPECT(*eff == *eff_back, "T5 efficient-storage round trip compares equal");
            EXPECT(!(*eff_back == *back), "T5 models with and without core matrix differ");
            EXPECT(std::fabs((*eff_back)(x, x) - (*gp)(x, x)) < 1e-9, "T5 core matrix is rebuilt on demand");
            EXPECT(!(*eff == *eff_back), "T5 after the lazy rebuild the core-matrix sizes differ");
            // float instantiation: cast at the boundary
            typedef GaussianProcess<float> GPf;
            GPf::Pointer gf = SinModel<float>(1.1f, 15, 2 * M_PI);
            gf->SetSigma(0.05f);
            gf->Initialize();
            gf->Save("/tmp/gprb200_cxx_io_f");
            GPf::Pointer gf_back(new GPf(Kernel<float>::Pointer(new WhiteKernel<float>(1))));
            gf_back->Load("/tmp/gprb200_cxx_io_f");
            EXPECT(*gf == *gf_back, "T5 float Save/Load round trip");
            EXPECT(std::fabs(gf->Predict(Vec1<float>(1.234f))[0] - (float)gp->Predict(x)[0]) < 1e-5, "T5 float model agrees with double");
        }
        // ---- T6 -----------------------------------------------------------------------------------------
        {
            typedef GaussianLogLikelihood<double> LL;
            Kernel<double>::Pointer k(new SumKernel<double>(Kernel<double>::Pointer(new GaussianKernel<double>(1.4, 1.2)),
                                                            Kernel<double>::Pointer(new PeriodicKernel<double>(0.7, 1.1, 0.9))));
            GP::Pointer gp(new GP(k));
            gp->SetSigma(0.3);
            for (unsigned i = 0; i < 40; ++i) {
                const double x = -2.0 + 0.1 * i;
                gp->AddSample(Vec1<double>(x), Vec1<double>(x * x + 0.3 * std::sin(5 * x)));
            }
            LL ll;
            const LL::ValueDerivativePair vd = ll.GetValueAndParameterDerivatives(gp);
            const Vec v0 = ll(gp);
            EXPECT(std::fabs(v0[0] - vd.first[0]) < 1e-10 * std::fabs(v0[0]), "T6 value from both entry points agrees");
            const Vec g = ll.GetParameterDerivatives(gp);
            const Kernel<double>::ParameterVectorType p0 = k->GetParameters();
            double worst = 0;
            for (size_t i = 0; i < p0.size(); ++i) {
                Kernel<double>::ParameterVectorType p = p0;
                const double h = 1e-5;
                p[i] = p0[i] + h;
                k->SetParameters(p);
                const double up = ll(gp)[0];
                p[i] = p0[i] - h;
                k->SetParameters(p);
                const double dn = ll(gp)[0];
                k->SetParameters(p0);
                worst = std::max(worst, std::fabs((up - dn) / (2 * h) - g[i]) / std::max(1.0, std::fabs(g[i])));
                EXPECT(std::fabs(g[i] - vd.second[i]) <= 1e-9 * std::max(1.0, std::fabs(g[i])), "T6 both gradient entry points agree");
            }
            std::cout << "T6 worst gradient error vs central differences = " << worst << std::endl;
            EXPECT(worst < 1e-5, "T6 analytic log-likelihood gradient matches finite differences");
            const LL::ValueJacobianPair vj = ll.GetValueAndJacobian(gp);
            EXPECT(vj.second.rows() == 1 && vj.second.cols() == (long)p0.size(), "T6 Jacobian shape M x P");
            EXPECT(std::fabs(vj.second(0, 0) - g[0]) <= 1e-8 * std::max(1.0, std::fabs(g[0])), "T6 Jacobian row equals gradient for M = 1");
            gp->SetDeterminantMode(GP::TrueLogDeterminant);
            EXPECT(std::fabs(ll(gp)[0] - v0[0]) < 1e-9 * std::fabs(v0[0]), "T6 clamp is inactive inside double range");
        }
        // ---- T7 -----------------------------------------------------------------------------------------
        {
            GP::Pointer gp(new GP(Kernel<double>::Pointer(new RationalQuadraticKernel<double>(1.0, 1.5, 2.0))));
            gp->SetSigma(0.2);
            for (unsigned i = 0; i < 130; ++i) {
                Vec x(3), y(5);
                for (int k = 0; k < 3; ++k) x[k] = std::sin(0.7 * i + k);
                for (int k = 0; k < 5; ++k) y[k] = std::cos(x[0] * (k + 1)) + x[1] * x[2];
                gp->AddSample(x, y);
            }
            const unsigned T = 11;
            Mat Xs(T, 3), mean;
            Vec ci;
            for (unsigned t = 0; t < T; ++t)
                for (int k = 0; k < 3; ++k) Xs(t, k) = std::cos(1.3 * t + k);
            gp->PredictBatch(Xs, mean, &ci);
            double worst_mean = 0, worst_ci = 0;
            for (unsigned t = 0; t < T; ++t) {
                Vec x(3);
                for (int k = 0; k < 3; ++k) x[k] = Xs(t, k);
                const Vec y = gp->Predict(x);
                for (int k = 0; k < 5; ++k) worst_mean = std::max(worst_mean, std::fabs(y[k] - mean(t, k)));
                worst_ci = std::max(worst_ci, std::fabs(gp->GetCredibleInterval(x) - ci[t]));
            }
            EXPECT(mean.rows() == T && mean.cols() == 5 && ci.rows() == T, "T7 batch shapes");
            EXPECT(worst_mean < 1e-12 && worst_ci < 1e-9, "T7 PredictBatch equals the one-point calls");
        }
        // ---- T8: label matrix >= 64 MB: Initialize takes the pipelined path (gprb200_learn_rows) -----------------
        {
            const unsigned N = 96, M = 100003;     // 76.8 MB of labels
            GP::Pointer big(new GP(Kernel<double>::Pointer(new GaussianKernel<double>(1.3, 1.0))));
            GP::Pointer one(new GP(Kernel<double>::Pointer(new GaussianKernel<double>(1.3, 1.0))));
            big->SetSigma(0.4);
            one->SetSigma(0.4);
            const unsigned probe = 77777;
            for (unsigned i = 0; i < N; ++i) {
                Vec x(2), y(M), y1(1);
                x[0] = std::sin(0.37 * i);
                x[1] = std::cos(0.11 * i * i);
                for (unsigned k = 0; k < M; ++k) y[k] = std::sin(x[0] * (1 + k % 13)) + 1e-3 * (k % 101) * x[1];
                y1[0] = y[probe];
                big->AddSample(x, y);
                one->AddSample(x, y1);
            }
            Vec xs(2);
            xs[0] = 0.21;
            xs[1] = -0.4;
            const Vec yb = big->Predict(xs);
            const Vec y1 = one->Predict(xs);
            EXPECT(yb.rows() == (long)M && std::fabs(yb[probe] - y1[0]) < 1e-11 * (1 + std::fabs(y1[0])),
                   "T8 pipelined Initialize: column of the 100003-column model equals the single-column model");
            big->Save("/tmp/gprb200_T8");
            GP::Pointer back(new GP(Kernel<double>::Pointer(new WhiteKernel<double>(1.0))));
            back->Load("/tmp/gprb200_T8");
            EXPECT(std::fabs(back->Predict(xs)[probe] - yb[probe]) < 1e-13, "T8 Save / Load of the large model");
        }
        // ---- T9: SetDistributed (world 1: the block-cyclic library's whole code path on one GPU, no exchange) ----------
        {
            const unsigned N = 1500;
            // (GetKernel consumes its argument, as the reference's does: one string per call)
            std::string ks = "SumKernel(GaussianKernel(1.3,1,),WhiteKernel(0.2,))", ks2 = ks, kg = "GaussianKernel(1.3,1,)";
            GP::Pointer dist(new GP(KernelFactory<double>::GetKernel(ks)));
            GP::Pointer plain(new GP(KernelFactory<double>::GetKernel(ks2)));
            dist->SetSigma(0.5);
            plain->SetSigma(0.5);
            for (unsigned i = 0; i < N; ++i) {
                Vec x(3), y(2);
                x[0] = std::sin(0.37 * i);
                x[1] = std::cos(0.11 * i * i);
                x[2] = std::sin(0.05 * i) * std::cos(0.3 * i);
                y[0] = std::sin(2 * x[0]) + 0.3 * x[1];
                y[1] = x[2] * x[0];
                dist->AddSample(x, y);
                plain->AddSample(x, y);
            }
            dist->SetDistributed(0, 1, nullptr);
            dist->Initialize();
            plain->Initialize();
            Vec xs(3);
            xs[0] = 0.2;
            xs[1] = -0.1;
            xs[2] = 0.05;
            const Vec a = dist->Predict(xs), b = plain->Predict(xs);
            EXPECT((a - b).norm() <= 1e-11 * (1 + b.norm()), "T9 SetDistributed: Predict equals the single-GPU model");
            EXPECT(std::fabs(dist->GetCredibleInterval(xs) - plain->GetCredibleInterval(xs)) <= 1e-10,
                   "T9 SetDistributed: credible interval through the solve on the grid");
            {   // batch of test points: T right-hand sides of one distributed solve
                Mat Xb(5, 3), mb, mp;
                for (int t = 0; t < 5; ++t) {
                    Xb(t, 0) = 0.1 * t - 0.2;
                    Xb(t, 1) = 0.3 - 0.05 * t;
                    Xb(t, 2) = 0.02 * t * t;
                }
                Vec cb, cp;
                dist->PredictBatch(Xb, mb, &cb);
                plain->PredictBatch(Xb, mp, &cp);
                EXPECT((mb - mp).norm() <= 1e-11 * (1 + mp.norm()) && (cb - cp).norm() <= 1e-9,
                       "T9 SetDistributed: PredictBatch with intervals (one solve with T right-hand sides on the grid)");
            }
            bool threw = false;
            try {
                GP::Pointer bad(new GP(KernelFactory<double>::GetKernel(kg)));
                bad->SetSigma(0);
                Vec x(1), y(1);
                x[0] = 0.5;
                y[0] = 1;
                for (int i = 0; i < 300; ++i) bad->AddSample(x, y);     // 300 identical samples, no noise: singular
                bad->SetDistributed(0, 1, nullptr);
                bad->Initialize();
            } catch (std::string& s) {
                threw = s.find("positive definite") != std::string::npos;
            }
            EXPECT(threw, "T9 SetDistributed: a singular kernel matrix throws the INFO message");
        }
    } catch (std::string& s) {
        std::cout << "[failed] exception: " << s << std::endl;
        return -1;
    }
    if (failures == 0) std::cout << "gp_device_test: [passed]" << std::endl;
    return failures == 0 ? 0 : -1;
}
