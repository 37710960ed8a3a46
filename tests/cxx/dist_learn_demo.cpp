// ONE GaussianProcess over several GPUs from plain C++ (no Python, no torch, no MPI): what a host application such as the
// reference's apps/GaussianProcessLearn.cpp:151 (gp->Initialize()) does to use the distributed Learn.
//   dist_learn_demo [world = 2] [N = 6000]
// The parent makes the NCCL id (GaussianProcess::GetDistributedId) and forks one child per GPU BEFORE touching CUDA; each
// child builds the same model, calls SetDevice(rank) + SetDistributed(rank, world, id) + Initialize() and predicts
// locally; rank 0 also trains a single-GPU model and compares. Exit code 0 and "[passed]" lines on success.
#include <sys/wait.h>
#include <unistd.h>

#include <cmath>
#include <cstdlib>
#include <iostream>
#include <string>
#include <vector>

#include "GaussianProcess.h"
#include "KernelFactory.h"

typedef gpr::GaussianProcess<double> GP;
typedef GP::VectorType Vec;

static GP::Pointer make_model(unsigned N) {
    std::string ks = "SumKernel(GaussianKernel(1.3,1,),WhiteKernel(0.2,))";
    GP::Pointer gp(new GP(gpr::KernelFactory<double>::GetKernel(ks)));
    gp->SetSigma(0.5);
    for (unsigned i = 0; i < N; ++i) {
        Vec x(3), y(2);
        x[0] = std::sin(0.37 * i);
        x[1] = std::cos(0.11 * i * i);
        x[2] = std::sin(0.05 * i) * std::cos(0.3 * i);
        y[0] = std::sin(2 * x[0]) + 0.3 * x[1];
        y[1] = x[2] * x[0];
        gp->AddSample(x, y);
    }
    return gp;
}

static int run_rank(int rank, int world, const std::vector<unsigned char>& id, unsigned N) {
    try {
        GP::Pointer gp = make_model(N);
        gp->SetDevice(rank);
        gp->SetDistributed(rank, world, id.data());
        gp->Initialize();
        Vec xs(3);
        xs[0] = 0.2;
        xs[1] = -0.1;
        xs[2] = 0.05;
        const Vec a = gp->Predict(xs);
        const double ci = gp->GetCredibleInterval(xs);      // collective: every rank asks
        if (rank == 0) {
            GP::Pointer plain = make_model(N);
            plain->SetDevice(0);
            const Vec b = plain->Predict(xs);
            const double ci1 = plain->GetCredibleInterval(xs);
            const bool ok = (a - b).norm() <= 1e-10 * (1 + b.norm()) && std::fabs(ci - ci1) <= 1e-9;
            std::cout << (ok ? "[passed]" : "[failed]") << " " << world << " ranks, N = " << N
                      << ": Predict and GetCredibleInterval after the distributed Initialize equal the single-GPU model (|diff| = "
                      << (a - b).norm() << ", " << std::fabs(ci - ci1) << ")" << std::endl;
            return ok ? 0 : 1;
        }
        return 0;
    } catch (std::string& s) {
        std::cout << "[failed] rank " << rank << ": " << s << std::endl;
        return 1;
    }
}

int main(int argc, char** argv) {
    const int world = argc > 1 ? std::atoi(argv[1]) : 2;
    const unsigned N = argc > 2 ? (unsigned)std::atoi(argv[2]) : 6000;
    std::vector<unsigned char> id;
    try {
        id = GP::GetDistributedId();    // starts NCCL's bootstrap listener in THIS process; no CUDA call yet
    } catch (std::string& s) {
        std::cout << "[failed] " << s << std::endl;
        return 1;
    }
    std::vector<pid_t> kids;
    for (int r = 0; r < world; ++r) {
        const pid_t p = fork();
        if (p == 0) _exit(run_rank(r, world, id, N));
        kids.push_back(p);
    }
    int bad = 0;
    for (pid_t p : kids) {
        int st = 0;
        waitpid(p, &st, 0);
        if (!WIFEXITED(st) || WEXITSTATUS(st) != 0) ++bad;
    }
    if (!bad) std::cout << "dist_learn_demo: [passed]" << std::endl;
    return bad ? 1 : 0;
}
