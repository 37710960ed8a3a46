// CPU-side checks of the C++ host API (no GPU needed): kernel arithmetic and parameter-derivative order
// (known answers of reference tests/KernelDerivativeTest.cpp:40-534: analytic gradient vs central differences
// at x=(0.1,0.5), y=(-0.1,0.8)), kernel string round trips (tests/HighlyGeneralKernelTest.cpp:73-150), MatrixIO
// (tests/IOTest.cpp:45-90), error strings, lowering to the device program, and the loud failure without a GPU.
#include <cmath>
#include <cstdio>
#include <iostream>
#include <string>
#include <vector>

#include "GaussianProcess.h"
#include "Kernel.h"
#include "KernelFactory.h"
#include "KernelUtils.h"
#include "MatrixIO.h"

using namespace gpr;
typedef Kernel<double> K;
typedef K::Pointer KP;
typedef K::VectorType Vec;

static int failures = 0;
#define EXPECT(cond, what)                                             \
    do {                                                               \
        if (!(cond)) {                                                 \
            ++failures;                                                \
            std::cout << "[failed] " << what << " (" #cond ")\n";      \
        }                                                              \
    } while (0)

static Vec V2(double a, double b) {
    Vec v(2);
    v[0] = a;
    v[1] = b;
    return v;
}

// analytic GetDerivative vs central differences over every parameter
static double GradientError(KP k, const Vec& x, const Vec& y) {
    const K::ParameterVectorType p0 = k->GetParameters();
    const Vec g = k->GetDerivative(x, y);
    double worst = 0;
    for (size_t i = 0; i < p0.size(); ++i) {
        const double h = 1e-6 * std::max(1.0, std::fabs(p0[i]));
        K::ParameterVectorType p = p0;
        p[i] = p0[i] + h;
        k->SetParameters(p);
        const double up = (*k)(x, y);
        p[i] = p0[i] - h;
        k->SetParameters(p);
        const double dn = (*k)(x, y);
        k->SetParameters(p0);
        worst = std::max(worst, std::fabs((up - dn) / (2 * h) - g[i]));
    }
    return worst;
}

int main() {
    const Vec x = V2(0.1, 0.5), y = V2(-0.1, 0.8);
    // ---- leaf values against closed forms -------------------------------------------------------------
    {
        const double r2 = 0.04 + 0.09;
        KP g(new GaussianKernel<double>(1.7, 2.5));
        EXPECT(std::fabs((*g)(x, y) - 2.5 * 2.5 * std::exp(-0.5 * r2 / (1.7 * 1.7))) < 1e-15, "Gaussian value");
        KP ge(new GaussianExpKernel<double>(0.3, -0.2));
        EXPECT(std::fabs((*ge)(x, y) - std::exp(-0.4) * std::exp(-0.5 * r2 / std::exp(0.6))) < 1e-15, "GaussianExp value");
        KP w(new WhiteKernel<double>(3.0));
        EXPECT((*w)(x, y) == 0.0 && (*w)(x, x) == 9.0, "White value fires on coincidence");
        KP rq(new RationalQuadraticKernel<double>(1.2, 0.9, 2.0));
        EXPECT(std::fabs((*rq)(x, y) - 1.44 * std::pow(1 + 0.5 * r2 / (0.81 * 2.0), -2.0)) < 1e-15, "RQ value");
        KP pk(new PeriodicKernel<double>(1.1, 2.0, 0.7));
        const double s = std::pow(std::sin(2.0 * 0.2), 2) + std::pow(std::sin(2.0 * -0.3), 2);
        EXPECT(std::fabs((*pk)(x, y) - 1.21 * std::exp(-0.5 * s / 0.49)) < 1e-15, "Periodic value");
    }
    // ---- gradients and their ORDER ----------------------------------------------------------------------
    {
        for (double sigma : {0.5, 1.0, 3.0})
            for (double scale : {0.5, 2.0}) {
                EXPECT(GradientError(KP(new GaussianKernel<double>(sigma, scale)), x, y) < 1e-8, "Gaussian gradient");
                EXPECT(GradientError(KP(new GaussianExpKernel<double>(std::log(sigma), std::log(scale))), x, y) < 1e-8,
                       "GaussianExp gradient");
                EXPECT(GradientError(KP(new RationalQuadraticKernel<double>(scale, sigma, 1.5)), x, y) < 1e-8, "RQ gradient");
                EXPECT(GradientError(KP(new PeriodicKernel<double>(scale, 1.3, sigma)), x, y) < 1e-8, "Periodic gradient");
            }
        EXPECT(GradientError(KP(new WhiteKernel<double>(2.0)), x, x) < 1e-8, "White gradient");
        KP g(new GaussianKernel<double>(1.5, 2.0)), p(new PeriodicKernel<double>(0.8, 1.2, 0.6));
        KP sum(new SumKernel<double>(g, p)), prod(new ProductKernel<double>(g, p));
        EXPECT(GradientError(sum, x, y) < 1e-8, "Sum gradient");
        EXPECT(GradientError(prod, x, y) < 1e-8, "Product gradient");
        // [g sigma, g scale, p scale, p b, p sigma]: k1's block first (reference tests :405-409)
        const Vec ds = sum->GetDerivative(x, y), dg = g->GetDerivative(x, y), dp = p->GetDerivative(x, y);
        EXPECT(ds.rows() == 5 && ds[0] == dg[0] && ds[1] == dg[1] && ds[2] == dp[0] && ds[4] == dp[2], "Sum gradient order");
        const Vec dq = prod->GetDerivative(x, y);
        EXPECT(std::fabs(dq[0] - dg[0] * (*p)(x, y)) < 1e-16 && std::fabs(dq[3] - dp[1] * (*g)(x, y)) < 1e-16,
               "Product gradient order");
        const K::ParameterVectorType ps = sum->GetParameters();
        EXPECT(ps.size() == 5 && ps[0] == 1.5 && ps[1] == 2.0 && ps[2] == 0.8 && ps[3] == 1.2 && ps[4] == 0.6,
               "Sum parameter order");
    }
    // ---- strings: ToString -> GetKernel -> equal kernel, nested ---------------------------------------------
    {
        std::vector<double> p13 = {1.0, 2.0, 0.5, 3.0, 1.5, 0.7, 2.5, 1.2, 0.9, 2.0, 0.3, 0.4, 0.01};
        KP general = GetGeneralKernel<double>(p13);
        EXPECT(general->GetNumberOfParameters() == 13, "general kernel has 13 parameters");
        std::string s = general->ToString();
        const std::string keep = s;
        KP again = KernelFactory<double>::GetKernel(s);
        EXPECT(again->ToString() == keep, "nested kernel string round trip");
        EXPECT(*again == *general, "nested kernel equality after round trip");
        EXPECT(std::fabs((*again)(x, y) - (*general)(x, y)) == 0.0, "round-tripped kernel evaluates identically");
        again->SetParameters(general->GetParameters());
        EXPECT(*again == *general, "SetParameters(GetParameters()) keeps equality");
        std::string prod_cfg = "GaussianKernel(35, 30,)";  // the production config string
        KP prodk = KernelFactory<double>::GetKernel(prod_cfg);
        EXPECT(prodk->GetParameters()[0] == 35.0 && prodk->GetParameters()[1] == 30.0, "config kernel string with blanks");
        EXPECT(prodk->ToString() == "GaussianKernel(35,30,)", "leaf ToString form");
        std::vector<int32_t> ops;
        std::vector<double> par;
        general->Lower(ops, par);
        const std::vector<int32_t> want = {1, 1, 5, 7, 6, 4, 6, 1, 3, 6, 6};
        EXPECT(ops == want, "post-order device program of the general kernel");
        EXPECT(par.size() == 13 && par[0] == 2.0 && par[1] == 1.0 && par[12] == 0.01, "flat parameters in GetParameters() order");
        bool threw = false;
        try {
            std::string bad = "NoSuchKernel(1,)";
            KernelFactory<double>::GetKernel(bad);
        } catch (std::string& e) {
            threw = (e == "KernelFactory::GetKernel: failed to load kernel.");
        }
        EXPECT(threw, "unknown kernel name throws the reference's message");
    }
    // ---- constructor guards (reference tests/GaussianProcessTest.cpp:322-344) -----------------------------------
    {
        bool a = false, b = false, c = false;
        try { GaussianKernel<double> k(0.0, 1.0); } catch (std::string& e) { a = (e == "GaussianKernel: sigma has to be positive"); }
        try { GaussianKernel<double> k(1.0, 0.0); } catch (std::string& e) { b = (e == "GaussianKernel: scale has to be positive"); }
        try { PeriodicKernel<double> k(1.0, 0.0, 1.0); } catch (std::string& e) { c = true; }
        EXPECT(a && b && c, "zero sigma / scale / period throw");
        bool d = false;
        try {
            GaussianKernel<double> k(1.0, 1.0);
            k.SetParameters(K::ParameterVectorType{1.0});
        } catch (std::string& e) { d = (e == "GaussianKernel::SetParameters: wrong number of parameters."); }
        EXPECT(d, "wrong parameter count throws");
    }
    // ---- MatrixIO ----------------------------------------------------------------------------------------
    {
        typedef Eigen::Matrix<double, Eigen::Dynamic, Eigen::Dynamic, Eigen::RowMajor> MD;
        typedef Eigen::Matrix<float, Eigen::Dynamic, Eigen::Dynamic, Eigen::RowMajor> MF;
        EXPECT(MatrixIOTest(), "MatrixIOTest");
        MD a(7, 5);
        for (int i = 0; i < 7; ++i)
            for (int j = 0; j < 5; ++j) a(i, j) = std::sin(1.0 + i * 5 + j);
        WriteMatrix<MD>(a, "/tmp/gprb200_host_test_d.txt");
        MD b = ReadMatrix<MD>("/tmp/gprb200_host_test_d.txt");
        bool same = b.rows() == 7 && b.cols() == 5;
        for (int i = 0; same && i < 7; ++i)
            for (int j = 0; j < 5; ++j) same = same && a(i, j) == b(i, j);
        EXPECT(same, "double matrix round trip is exact");
        FILE* f = std::fopen("/tmp/gprb200_host_test_d.txt", "rb");
        char hdr[8] = {0};
        std::fread(hdr, 1, 4, f);
        std::fseek(f, 0, SEEK_END);
        const long bytes = std::ftell(f);
        std::fclose(f);
        EXPECT(std::string(hdr, 4) == "7 5\n" && bytes == 4 + 7 * 5 * 8, "file = 'rows cols\\n' + raw row-major scalars");
        MF c(3, 2);
        for (int i = 0; i < 6; ++i) c.data()[i] = 0.1f * i;
        WriteMatrix<MF>(c, "/tmp/gprb200_host_test_f.txt");
        MF d = ReadMatrix<MF>("/tmp/gprb200_host_test_f.txt");
        EXPECT(d.rows() == 3 && d.cols() == 2 && d(2, 1) == c(2, 1), "float matrix round trip");
        MD e(0, 0);
        WriteMatrix<MD>(e, "/tmp/gprb200_host_test_e.txt");
        MD e2 = ReadMatrix<MD>("/tmp/gprb200_host_test_e.txt");
        EXPECT(e2.rows() == 0 && e2.cols() == 0, "empty (efficient-storage) core matrix round trip");
    }
    // ---- GaussianProcess host logic; and no silent CPU fallback ---------------------------------------------------
    {
        typedef GaussianProcess<double> GP;
        GP::Pointer gp(new GP(KP(new GaussianKernel<double>(1.0, 1.0))));
        bool a = false, b = false, c = false;
        try { gp->Initialize(); } catch (std::string& e) { a = (e.find("no input samples") != std::string::npos); }
        Vec x1(1), y1(1), x2 = V2(1, 2);
        x1[0] = 0.5;
        y1[0] = 1.0;
        gp->AddSample(x1, y1);
        try { gp->AddSample(x2, y1); } catch (std::string& e) { b = (e.find("does not correspond to the input dimension (1)") != std::string::npos); }
        try { gp->Save("/tmp/gprb200_host_test_gp"); } catch (std::string& e) { c = (e == "GaussianProcess::Save: gaussian process is not initialized."); }
        EXPECT(a && b && c, "Initialize/AddSample/Save error strings");
        EXPECT(gp->GetNumberOfSamples() == 1 && gp->GetNumberOfInputDimensions() == 1, "sample bookkeeping");
        gp->SetSigma(0.25);
        EXPECT(gp->GetSigmaSquared() == 0.0625, "sigma accessors");
        if (std::getenv("GPRB200_EXPECT_NO_GPU")) {
            bool loud = false;
            try { gp->Predict(x1); } catch (std::string& e) { loud = (e.find("no CPU fallback") != std::string::npos); }
            EXPECT(loud, "without a GPU Predict throws instead of computing on the CPU");
        }
    }
    if (failures == 0) std::cout << "host_test: [passed]" << std::endl;
    return failures == 0 ? 0 : -1;
}
