// Times the loop body of the reference's gpPredict (apps/GaussianProcessPredict.cpp:185-193) through the host C++ API
// (libgplib.so over libgprb200.so): for each test point ONE gp->Predict(x) and ONE gp->GetCredibleInterval(x).
//   predict_loop_bench N d M T   ->  one JSON line on stdout (needs a B200)
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <memory>
#include <string>

#include "GaussianProcess.h"
#include "Kernel.h"

using namespace gpr;

int main(int argc, char** argv) {
    typedef GaussianProcess<double> GP;
    typedef GP::VectorType Vec;
    const long N = argc > 1 ? std::atol(argv[1]) : 16384, d = argc > 2 ? std::atol(argv[2]) : 8;
    const long M = argc > 3 ? std::atol(argv[3]) : 4, T = argc > 4 ? std::atol(argv[4]) : 200;
    try {
        GP::Pointer gp(new GP(Kernel<double>::Pointer(new GaussianKernel<double>(2.0, 1.0))));
        gp->SetSigma(1.0);
        unsigned long long s = 88172645463325252ull;
        auto rnd = [&]() { s ^= s << 13; s ^= s >> 7; s ^= s << 17; return (double)(s >> 11) / 9007199254740992.0 * 2.0 - 1.0; };
        for (long i = 0; i < N; ++i) {
            Vec x(d), y(M);
            for (long k = 0; k < d; ++k) x[k] = rnd();
            for (long k = 0; k < M; ++k) y[k] = std::sin(3.0 * x[0] + k) + 0.1 * rnd();
            gp->AddSample(x, y);
        }
        auto t0 = std::chrono::steady_clock::now();
        gp->Initialize();
        const double learn_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        Vec x(d);
        double acc = 0.0, first_ms = 0.0;
        for (long t = 0; t < 8; ++t) {   // warm-up: includes the one-off inverse-factor build of the interval path
            for (long k = 0; k < d; ++k) x[k] = rnd();
            auto w0 = std::chrono::steady_clock::now();
            acc += gp->Predict(x)[0] + gp->GetCredibleInterval(x);
            if (t < 3) first_ms += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - w0).count();
        }
        t0 = std::chrono::steady_clock::now();
        double tp = 0.0;
        for (long t = 0; t < T; ++t) {
            for (long k = 0; k < d; ++k) x[k] = rnd();
            auto p0 = std::chrono::steady_clock::now();
            const Vec y = gp->Predict(x);
            tp += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - p0).count();
            acc += y[0] + gp->GetCredibleInterval(x);
        }
        const double loop_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        std::printf("{\"N\": %ld, \"d\": %ld, \"M\": %ld, \"T\": %ld, \"learn_ms\": %.3f, \"ms_per_point\": %.4f, "
                    "\"predict_ms_per_point\": %.4f, \"interval_ms_per_point\": %.4f, \"first_three_points_ms\": %.3f, "
                    "\"checksum\": %.6f}\n",
                    N, d, M, T, learn_ms, loop_ms / T, tp / T, (loop_ms - tp) / T, first_ms, acc);
    } catch (std::string& e) {
        std::fprintf(stderr, "predict_loop_bench: %s\n", e.c_str());
        return 1;
    }
    return 0;
}
