// gpr::GaussianProcess<T> over the C-ABI of include/gprb200.h (libgprb200.so, sm_100a CUDA).
// Control flow follows the reference's lib/GaussianProcess.cpp (cited per method); the numerical
// work is on the GPU. No CPU implementation of the path exists here: without the CUDA library and
// a B200 every entry point throws.
#include <cmath>
#include <cstdlib>
#include <fstream>
#include <iomanip>
#include <limits>
#include <mutex>
#include <sstream>
#include <type_traits>
#include <dlfcn.h>
#include <sys/stat.h>

#include "GaussianProcess.h"
#include "KernelFactory.h"
#include "MatrixIO.h"
#include "gprb200.h"
#include "gprb200_dist.h"

namespace gpr {

namespace detail {

struct DeviceModel {
    gprb200_handle h = nullptr;
    std::vector<int32_t> sent_ops;
    std::vector<double> sent_params;
    double sent_sigma = std::numeric_limits<double>::quiet_NaN();

    explicit DeviceModel(int device) {
        const int rc = gprb200_create(&h, device);
        if (rc != 0) throw std::string("GaussianProcess: ") + gprb200_last_error(nullptr);
    }
    ~DeviceModel() {
        if (h) gprb200_destroy(h);
    }
    void Check(int rc) const {
        if (rc != 0) throw std::string(gprb200_last_error(h));
    }
};

// libgprb200_dist.so, opened on first use (it needs NCCL; libgplib.so must not), and this rank's handle in it
struct DistLib {
    void* so = nullptr;
    decltype(&gprb200_dist_get_unique_id) get_unique_id = nullptr;
    decltype(&gprb200_dist_create) create = nullptr;
    decltype(&gprb200_dist_destroy) destroy = nullptr;
    decltype(&gprb200_dist_last_error) last_error = nullptr;
    decltype(&gprb200_dist_set_kernel) set_kernel = nullptr;
    decltype(&gprb200_dist_learn) learn = nullptr;
    decltype(&gprb200_dist_solve) solve = nullptr;

    static DistLib& Get() {
        static DistLib lib;
        static std::mutex guard;    // ranks may be threads of one process
        std::lock_guard<std::mutex> lock(guard);
        if (!lib.so) {
            lib.so = ::dlopen("libgprb200_dist.so", RTLD_NOW | RTLD_GLOBAL);   // found through libgplib.so's RUNPATH
            if (!lib.so) throw std::string("GaussianProcess::SetDistributed: cannot open libgprb200_dist.so: ") + ::dlerror();
            auto sym = [&](const char* name) {
                void* p = ::dlsym(lib.so, name);
                if (!p) throw std::string("GaussianProcess::SetDistributed: libgprb200_dist.so lacks ") + name;
                return p;
            };
            lib.get_unique_id = reinterpret_cast<decltype(lib.get_unique_id)>(sym("gprb200_dist_get_unique_id"));
            lib.create = reinterpret_cast<decltype(lib.create)>(sym("gprb200_dist_create"));
            lib.destroy = reinterpret_cast<decltype(lib.destroy)>(sym("gprb200_dist_destroy"));
            lib.last_error = reinterpret_cast<decltype(lib.last_error)>(sym("gprb200_dist_last_error"));
            lib.set_kernel = reinterpret_cast<decltype(lib.set_kernel)>(sym("gprb200_dist_set_kernel"));
            lib.learn = reinterpret_cast<decltype(lib.learn)>(sym("gprb200_dist_learn"));
            lib.solve = reinterpret_cast<decltype(lib.solve)>(sym("gprb200_dist_solve"));
        }
        return lib;
    }
};

struct DistModel {
    gprb200_dist_handle d = nullptr;
    double logdet = 0.0;
    DistModel(int device, int rank, int world, const void* id, int Pr, int Pc) {
        DistLib& lib = DistLib::Get();
        if (lib.create(&d, device, rank, world, id, Pr, Pc) != 0)
            throw std::string("GaussianProcess::SetDistributed: ") + lib.last_error(nullptr);
    }
    ~DistModel() {
        if (d) DistLib::Get().destroy(d);
    }
};

}  // namespace detail

namespace {

bool IsRegularFile(const std::string& fn) {
    struct stat st;
    return ::stat(fn.c_str(), &st) == 0 && !S_ISDIR(st.st_mode);
}

// T <-> double staging at the C-ABI (the float instantiation casts at the boundary, like the
// reference does around LAPACK: include/LAPACKUtils.h:87,96)
template <class T>
struct Staging {
    std::vector<double> buf;
    const double* In(const T* src, size_t n) {
        buf.assign(src, src + n);
        return buf.data();
    }
    double* Out(T*, size_t n) {
        buf.resize(n);
        return buf.data();
    }
    void Finish(T* dst, size_t n) {
        for (size_t i = 0; i < n; ++i) dst[i] = static_cast<T>(buf[i]);
    }
};
template <>
struct Staging<double> {
    const double* In(const double* src, size_t) { return src; }
    double* Out(double* dst, size_t) { return dst; }
    void Finish(double*, size_t) {}
};

}  // namespace

// ---- construction -----------------------------------------------------------------------------------
template <class TScalarType>
GaussianProcess<TScalarType>::GaussianProcess(KernelTypePointer kernel)
    : m_Kernel(kernel),
      m_Sigma(0),
      m_Initialized(false),
      m_InputDimension(0),
      m_OutputDimension(0),
      m_InvMethod(FullPivotLU),
      m_EfficientStorage(false),
      debug(false),
      m_DeterminantMode(ReferenceClamp),
      m_CoreSize(0),
      m_TrainDirty(true),
      m_AlphaMirrored(false),
      m_CoreMirrored(false),
      m_DeviceIndex(-1) {}

template <class TScalarType>
GaussianProcess<TScalarType>::~GaussianProcess() {
    if (debug) std::cout << "GaussianProcess::~GaussianProcess() destruct object" << std::endl;
}

template <class TScalarType>
void GaussianProcess<TScalarType>::SetDevice(int device) {
    if (m_Device && device != m_DeviceIndex) {
        m_Device.reset();
        m_TrainDirty = true;
        m_Initialized = false;
    }
    m_DeviceIndex = device;
}

template <class TScalarType>
detail::DeviceModel& GaussianProcess<TScalarType>::Device() const {
    if (!m_Device) {
        int device = m_DeviceIndex;
        if (device < 0) {
            const char* env = std::getenv("GPRB200_DEVICE");
            device = env ? std::atoi(env) : 0;
        }
        m_Device.reset(new detail::DeviceModel(device));
    }
    return *m_Device;
}

// kernel program + training data + sigma -> device, sending only what changed
template <class TScalarType>
void GaussianProcess<TScalarType>::SyncModelToDevice() const {
    detail::DeviceModel& dev = Device();
    std::vector<int32_t> ops;
    std::vector<double> params;
    m_Kernel->Lower(ops, params);
    if (ops != dev.sent_ops) {
        dev.Check(gprb200_set_kernel(dev.h, ops.data(), (int)ops.size(), params.data(), (int)params.size()));
    } else if (params != dev.sent_params) {
        dev.Check(gprb200_set_kernel_params(dev.h, params.data(), (int)params.size()));
    }
    dev.sent_ops = ops;
    dev.sent_params = params;

    if (m_TrainDirty) {
        const size_t N = m_SampleVectors.size(), d = m_InputDimension, M = m_OutputDimension;
        if (N == 0) throw std::string("GaussianProcess: no input samples defined");
        if (std::is_same<TScalarType, double>::value) {
            // the vectors go to the device from where AddSample left them (one pointer per vector, labels streamed
            // through pinned staging): no N x M host gather (reference: ComputeLabelMatrixInternal, :621-639)
            std::vector<const double*> xr(N), yr(N);
            for (size_t i = 0; i < N; ++i) {
                xr[i] = reinterpret_cast<const double*>(m_SampleVectors[i].data());
                yr[i] = M ? reinterpret_cast<const double*>(m_LabelVectors[i].data()) : nullptr;
            }
            dev.Check(gprb200_set_train_rows(dev.h, xr.data(), (int64_t)N, (int)d, M ? yr.data() : nullptr, (int64_t)M,
                                             (double)m_Sigma));
        } else {
            std::vector<double> X(N * d), Y(N * M);
            for (size_t i = 0; i < N; ++i) {
                for (size_t k = 0; k < d; ++k) X[i * d + k] = m_SampleVectors[i][k];
                for (size_t k = 0; k < M; ++k) Y[i * M + k] = m_LabelVectors[i][k];
            }
            dev.Check(gprb200_set_train(dev.h, X.data(), (int64_t)N, (int)d, M ? Y.data() : nullptr, (int64_t)M,
                                        (double)m_Sigma));
        }
        dev.sent_sigma = (double)m_Sigma;
        m_TrainDirty = false;
    }
    if (dev.sent_sigma != (double)m_Sigma) {
        dev.Check(gprb200_set_sigma(dev.h, (double)m_Sigma));
        dev.sent_sigma = (double)m_Sigma;
    }
}

// ---- public API -------------------------------------------------------------------------------------
// reference lib/GaussianProcess.cpp:36-51
template <class TScalarType>
void GaussianProcess<TScalarType>::AddSample(const VectorType& x, const VectorType& y) {
    if (m_SampleVectors.empty()) m_InputDimension = x.size();
    if (m_LabelVectors.empty()) m_OutputDimension = y.size();
    CheckInputDimension(x, "GaussianProcess::AddSample: ");
    CheckOutputDimension(y, "GaussianProcess::AddSample: ");
    m_SampleVectors.push_back(x);
    m_LabelVectors.push_back(y);
    m_Initialized = false;
    m_TrainDirty = true;
}

// reference :118-130
template <class TScalarType>
void GaussianProcess<TScalarType>::Initialize() {
    if (m_Initialized) return;
    if (!(m_SampleVectors.size() > 0))
        throw std::string("GaussianProcess::Initialize: no input samples defined during initialization");
    if (!(m_LabelVectors.size() > 0))
        throw std::string("GaussianProcess::Initialize: no ouput labels defined during initialization");
    ComputeRegressionVectors();
    m_Initialized = true;
}

template <class TScalarType>
std::vector<unsigned char> GaussianProcess<TScalarType>::GetDistributedId() {
    std::vector<unsigned char> id(GPRB200_DIST_ID_BYTES);
    detail::DistLib& lib = detail::DistLib::Get();
    if (lib.get_unique_id(id.data()) != 0) throw std::string("GaussianProcess::GetDistributedId: ") + lib.last_error(nullptr);
    return id;
}

template <class TScalarType>
void GaussianProcess<TScalarType>::SetDistributed(int rank, int world, const void* nccl_id, int Pr, int Pc) {
    int device = m_DeviceIndex;
    if (device < 0) {
        const char* env = std::getenv("GPRB200_DEVICE");
        device = env ? std::atoi(env) : 0;
    }
    m_Dist.reset(new detail::DistModel(device, rank, world, nccl_id, Pr, Pc));
    m_Initialized = false;
}

// ComputeRegressionVectors over the GPUs of the grid (SetDistributed): alpha = (K + sigma^2 I)^-1 Y by the block-cyclic
// Cholesky of libgprb200_dist.so; this rank then holds alpha on the host and on its own GPU, installed like a loaded
// model (no local factor: a credible interval asked for later factors K on this GPU alone, if it fits).
template <class TScalarType>
void GaussianProcess<TScalarType>::ComputeRegressionVectorsDistributed() {
    detail::DistLib& lib = detail::DistLib::Get();
    const size_t N = m_SampleVectors.size(), d = m_InputDimension, M = m_OutputDimension;
    std::vector<int32_t> ops;
    std::vector<double> params;
    m_Kernel->Lower(ops, params);
    std::vector<double> X(N * d), Y(N * M), A(N * M);
    for (size_t i = 0; i < N; ++i) {
        for (size_t k = 0; k < d; ++k) X[i * d + k] = m_SampleVectors[i][k];
        for (size_t k = 0; k < M; ++k) Y[i * M + k] = m_LabelVectors[i][k];
    }
    if (lib.set_kernel(m_Dist->d, ops.data(), (int)ops.size(), params.data(), (int)params.size()) != 0)
        throw std::string(lib.last_error(m_Dist->d));
    const int rc = lib.learn(m_Dist->d, X.data(), (int64_t)N, (int)d, Y.data(), (int64_t)M, (double)m_Sigma, 0, A.data(),
                             &m_Dist->logdet);
    if (rc != 0) throw std::string(lib.last_error(m_Dist->d));
    m_RegressionVectors.resize(N, M);
    for (size_t i = 0; i < N * M; ++i) m_RegressionVectors.data()[i] = (TScalarType)A[i];
    detail::DeviceModel& dev = Device();
    dev.Check(gprb200_set_kernel(dev.h, ops.data(), (int)ops.size(), params.data(), (int)params.size()));
    dev.sent_ops = ops;
    dev.sent_params = params;
    dev.Check(gprb200_set_state(dev.h, X.data(), (int64_t)N, (int)d, Y.data(), A.data(), (int64_t)M, nullptr, (double)m_Sigma));
    dev.sent_sigma = (double)m_Sigma;
    m_TrainDirty = false;
    m_AlphaMirrored = true;
    m_CoreMirrored = false;
    m_CoreMatrix.resize(0, 0);
    m_CoreSize = m_EfficientStorage ? 0 : (unsigned)N;
}

// reference :642-672 — Gram + noise, factor, alpha; C is not formed here (see MirrorCoreMatrix)
template <class TScalarType>
void GaussianProcess<TScalarType>::ComputeRegressionVectors() {
    if (m_Dist) {
        ComputeRegressionVectorsDistributed();
        return;
    }
    const size_t N0 = m_SampleVectors.size(), M0 = m_OutputDimension;
    if (std::is_same<TScalarType, double>::value && m_TrainDirty && N0 * M0 * sizeof(double) >= (size_t(64) << 20)) {
        // large label matrix: upload and solve overlapped over column blocks, straight from the AddSample storage
        m_TrainDirty = false;    // SyncModelToDevice below only brings the kernel up to date
        try {
            SyncModelToDevice();
        } catch (...) {
            m_TrainDirty = true;
            throw;
        }
        detail::DeviceModel& dev = Device();
        std::vector<const double*> xr(N0), yr(N0);
        for (size_t i = 0; i < N0; ++i) {
            xr[i] = reinterpret_cast<const double*>(m_SampleVectors[i].data());
            yr[i] = reinterpret_cast<const double*>(m_LabelVectors[i].data());
        }
        m_TrainDirty = true;     // restored if the call throws
        dev.Check(gprb200_learn_rows(dev.h, xr.data(), (int64_t)N0, (int)m_InputDimension, yr.data(), (int64_t)M0,
                                     (double)m_Sigma, nullptr));
        dev.sent_sigma = (double)m_Sigma;
        m_TrainDirty = false;
    } else {
        SyncModelToDevice();
        detail::DeviceModel& dev = Device();
        dev.Check(gprb200_learn(dev.h, 0u));
    }
    m_AlphaMirrored = false;
    m_CoreMirrored = false;
    m_RegressionVectors.resize(0, 0);
    m_CoreMatrix.resize(0, 0);
    m_CoreSize = m_EfficientStorage ? 0 : m_SampleVectors.size();
    // small models keep the reference's eager host copy of the regression vectors
    if ((double)m_SampleVectors.size() * (double)m_OutputDimension <= 4194304.0) MirrorRegressionVectors();
}

// reference :55-61
template <class TScalarType>
typename GaussianProcess<TScalarType>::VectorType GaussianProcess<TScalarType>::Predict(const VectorType& x) {
    Initialize();
    CheckInputDimension(x, "GaussianProcess::Predict: ");
    SyncModelToDevice();
    detail::DeviceModel& dev = Device();
    VectorType y = VectorType::Zero(m_OutputDimension);
    Staging<TScalarType> sx, sy;
    const double* px = sx.In(x.data(), m_InputDimension);
    double* py = sy.Out(y.data(), m_OutputDimension);
    dev.Check(gprb200_predict(dev.h, px, 1, py, nullptr));
    sy.Finish(y.data(), m_OutputDimension);
    return y;
}

template <class TScalarType>
void GaussianProcess<TScalarType>::PredictBatch(const MatrixType& Xs, MatrixType& mean, VectorType* ci) {
    Initialize();
    if ((unsigned)Xs.cols() != m_InputDimension) {
        std::stringstream ss;
        ss << "GaussianProcess::PredictBatch: dimension of input vector (" << Xs.cols()
           << ") does not correspond to the input dimension (" << m_InputDimension << ").";
        throw ss.str();
    }
    SyncModelToDevice();
    detail::DeviceModel& dev = Device();
    const size_t T = Xs.rows();
    mean.resize(T, m_OutputDimension);
    if (ci) ci->resize(T);
    Staging<TScalarType> sx, sm, sc;
    const double* px = sx.In(Xs.data(), T * m_InputDimension);
    double* pm = sm.Out(mean.data(), T * m_OutputDimension);
    double* pc = ci ? sc.Out(ci->data(), T) : nullptr;
    if (ci && m_Dist) {
        // the factor lives on the GPU grid: the T kernel vectors are T right-hand sides of ONE distributed solve
        // (collective: every rank makes the same call), ci_t = 2 sqrt(max(0, k(x_t, x_t) - Kx_t^T (K + s^2 I)^-1 Kx_t))
        dev.Check(gprb200_predict(dev.h, px, (int64_t)T, pm, nullptr));
        detail::DistLib& lib = detail::DistLib::Get();
        const size_t N = m_SampleVectors.size();
        std::vector<double> Kx(T * N), Kt(N * T), S(N * T);
        dev.Check(gprb200_kernel_vector(dev.h, px, (int64_t)T, Kx.data()));
        for (size_t t = 0; t < T; ++t)
            for (size_t i = 0; i < N; ++i) Kt[i * T + t] = Kx[t * N + i];
        if (lib.solve(m_Dist->d, Kt.data(), (int64_t)T, S.data()) != 0) throw std::string(lib.last_error(m_Dist->d));
        for (size_t t = 0; t < T; ++t) {
            long double acc = 0;
            for (size_t i = 0; i < N; ++i) acc += (long double)Kt[i * T + t] * S[i * T + t];
            VectorType xt = Xs.row(t).transpose();
            const double c = (double)(*m_Kernel)(xt, xt) - (double)acc;
            pc[t] = c > 0 ? 2.0 * std::sqrt(c) : 0.0;
        }
    } else {
        dev.Check(gprb200_predict(dev.h, px, (int64_t)T, pm, pc));
    }
    sm.Finish(mean.data(), T * m_OutputDimension);
    if (ci) sc.Finish(ci->data(), T);
    if (ci && m_CoreSize == 0) m_CoreSize = m_SampleVectors.size();
}

// reference :65-81 (D = -Xd^T (Kx o alpha_i); no 1/sigma^2 factor, as in the reference)
template <class TScalarType>
typename GaussianProcess<TScalarType>::VectorType GaussianProcess<TScalarType>::PredictDerivative(const VectorType& x,
                                                                                                    MatrixType& D) {
    Initialize();
    CheckInputDimension(x, "GaussianProcess::PredictDerivative: ");
    SyncModelToDevice();
    detail::DeviceModel& dev = Device();
    VectorType y = VectorType::Zero(m_OutputDimension);
    D.resize(m_InputDimension, m_OutputDimension);
    Staging<TScalarType> sx, sy, sd;
    const double* px = sx.In(x.data(), m_InputDimension);
    double* py = sy.Out(y.data(), m_OutputDimension);
    double* pd = sd.Out(D.data(), (size_t)m_InputDimension * m_OutputDimension);
    dev.Check(gprb200_predict_derivative(dev.h, px, py, pd));
    sy.Finish(y.data(), m_OutputDimension);
    sd.Finish(D.data(), (size_t)m_InputDimension * m_OutputDimension);
    return y;
}

// reference :85-99 — k(x,y) - Kx^T C Ky; with a factor on the device this is k - (L^-1 Kx).(L^-1 Ky)
template <class TScalarType>
TScalarType GaussianProcess<TScalarType>::operator()(const VectorType& x, const VectorType& y) {
    Initialize();
    CheckInputDimension(x, "GaussianProcess::(): ");
    CheckInputDimension(y, "GaussianProcess::(): ");
    SyncModelToDevice();
    detail::DeviceModel& dev = Device();
    if (m_CoreSize == 0) m_CoreSize = m_SampleVectors.size();  // the reference rebuilds m_CoreMatrix here (:95-97)
    Staging<TScalarType> sx, sy;
    const double* px = sx.In(x.data(), m_InputDimension);
    const double* py = sy.In(y.data(), m_InputDimension);
    double cov = 0.0;
    if (m_Dist) {
        // the factor lives on the GPU grid, not on this GPU: gp(x, y) = k(x, y) - Kx^T (K + sigma^2 I)^-1 Ky with the solve
        // on the grid. COLLECTIVE: every rank makes the same call (each gets the value).
        detail::DistLib& lib = detail::DistLib::Get();
        const size_t N = m_SampleVectors.size();
        std::vector<double> Kx(N), Ky(N), S(N);
        dev.Check(gprb200_kernel_vector(dev.h, px, 1, Kx.data()));
        dev.Check(gprb200_kernel_vector(dev.h, py, 1, Ky.data()));
        if (lib.solve(m_Dist->d, Ky.data(), 1, S.data()) != 0) throw std::string(lib.last_error(m_Dist->d));
        long double acc = 0;
        for (size_t i = 0; i < N; ++i) acc += (long double)Kx[i] * S[i];
        return static_cast<TScalarType>((double)(*m_Kernel)(x, y) - (double)acc);
    }
    dev.Check(gprb200_posterior_cov(dev.h, px, py, 1, &cov));
    return static_cast<TScalarType>(cov);
}

// reference :103-114
template <class TScalarType>
TScalarType GaussianProcess<TScalarType>::GetCredibleInterval(const VectorType& x) {
    Initialize();
    CheckInputDimension(x, "GaussianProcess::GetCredibleIntervall: ");
    const TScalarType c = (*this)(x, x);
    if (debug && c < 0)
        std::cout << "GaussianProcess::GetCredibleIntervall: prediction is instable. gp(x,x) = " << c << "." << std::endl;
    return !(c > 0) ? TScalarType(0) : TScalarType(2 * std::sqrt(c));
}

template <class TScalarType>
void* GaussianProcess<TScalarType>::SyncedDeviceHandle() const {
    SyncModelToDevice();
    return Device().h;
}

template <class TScalarType>
void* GaussianProcess<TScalarType>::GetDeviceHandle() {
    Initialize();
    return Device().h;
}

template <class TScalarType>
std::vector<double> GaussianProcess<TScalarType>::GetPhaseMilliseconds() const {
    std::vector<double> ms(GPRB200_T_COUNT, 0.0);
    if (m_Device) gprb200_phase_ms(m_Device->h, ms.data());
    return ms;
}

// ---- host mirrors -----------------------------------------------------------------------------------
template <class TScalarType>
void GaussianProcess<TScalarType>::MirrorRegressionVectors() const {
    if (m_AlphaMirrored) return;
    detail::DeviceModel& dev = Device();
    const size_t N = m_SampleVectors.size(), M = m_OutputDimension;
    m_RegressionVectors.resize(N, M);
    Staging<TScalarType> st;
    double* p = st.Out(m_RegressionVectors.data(), N * M);
    dev.Check(gprb200_get_alpha(dev.h, p));
    st.Finish(m_RegressionVectors.data(), N * M);
    m_AlphaMirrored = true;
}

// C = inv(K + sigma^2 I) formed on the device from the Cholesky factor (InvertKernelMatrix,
// reference :531-618) the first time the host needs it
template <class TScalarType>
void GaussianProcess<TScalarType>::MirrorCoreMatrix() const {
    if (m_CoreMirrored) return;
    if (m_CoreSize == 0) {
        m_CoreMatrix.resize(0, 0);
        m_CoreMirrored = true;
        return;
    }
    detail::DeviceModel& dev = Device();
    const size_t N = m_SampleVectors.size();
    m_CoreMatrix.resize(N, N);
    Staging<TScalarType> st;
    double* p = st.Out(m_CoreMatrix.data(), N * N);
    dev.Check(gprb200_get_core(dev.h, p));
    st.Finish(m_CoreMatrix.data(), N * N);
    m_CoreMirrored = true;
}

// ---- Save / Load: the reference's five-file model (reference :134-268) ---------------------------------
template <class TScalarType>
void GaussianProcess<TScalarType>::Save(std::string prefix) {
    if (!m_Initialized) throw std::string("GaussianProcess::Save: gaussian process is not initialized.");
    if (debug) {
        std::cout << "GaussianProcess::Save: writing gaussian process: " << std::endl;
        const char* tags[] = {"-RegressionVectors.txt", "-CoreMatrix.txt", "-SampleVectors.txt", "-LabelVectors.txt",
                              "-ParameterFile.txt"};
        for (const char* t : tags) std::cout << "\t " << prefix + t << std::endl;
    }
    SyncModelToDevice();
    MirrorRegressionVectors();
    WriteMatrix<MatrixType>(m_RegressionVectors, prefix + "-RegressionVectors.txt");

    if (m_EfficientStorage) {  // the reference drops the core matrix here for good (:152)
        m_CoreSize = 0;
        m_CoreMirrored = false;
    }
    MirrorCoreMatrix();
    WriteMatrix<MatrixType>(m_CoreMatrix, prefix + "-CoreMatrix.txt");

    const size_t N = m_SampleVectors.size();
    MatrixType X = MatrixType::Zero(m_InputDimension, N);  // one sample per COLUMN (:156-160)
    MatrixType Y = MatrixType::Zero(m_OutputDimension, N);
    for (size_t i = 0; i < N; ++i) {
        for (unsigned k = 0; k < m_InputDimension; ++k) X(k, i) = m_SampleVectors[i][k];
        for (unsigned k = 0; k < m_OutputDimension; ++k) Y(k, i) = m_LabelVectors[i][k];
    }
    WriteMatrix<MatrixType>(X, prefix + "-SampleVectors.txt");
    WriteMatrix<MatrixType>(Y, prefix + "-LabelVectors.txt");

    // sigma goes out at the stream's default precision, the kernel string at full precision (:174-177)
    std::ofstream pf((prefix + "-ParameterFile.txt").c_str());
    pf << m_Sigma << " " << m_InputDimension << " " << m_OutputDimension << " " << m_EfficientStorage << " " << debug
       << " ";
    pf << std::setprecision(std::numeric_limits<TScalarType>::digits10 + 1);
    pf << m_Kernel->ToString();
}

template <class TScalarType>
void GaussianProcess<TScalarType>::Load(std::string prefix) {
    const char* tags[] = {"-RegressionVectors.txt", "-CoreMatrix.txt", "-SampleVectors.txt", "-LabelVectors.txt",
                          "-ParameterFile.txt"};
    if (debug) {
        std::cout << "GaussianProcess::Load: loading gaussian process: " << std::endl;
        for (const char* t : tags) std::cout << "\t " << prefix + t << std::endl;
    }
    // same order of checks and reads as the reference
    MatrixType loaded[4];
    for (int f = 0; f < 4; ++f) {
        const std::string fn = prefix + tags[f];
        if (!IsRegularFile(fn)) throw std::string("GaussianProcess::Load: " + fn + " does not exist or is a directory.");
        loaded[f] = ReadMatrix<MatrixType>(fn);
    }
    const std::string pf_name = prefix + tags[4];
    if (!IsRegularFile(pf_name))
        throw std::string("GaussianProcess::Load: " + pf_name + " does not exist or is a directory.");

    m_RegressionVectors = loaded[0];
    m_CoreMatrix = loaded[1];
    const MatrixType& X = loaded[2];
    const MatrixType& Y = loaded[3];
    m_SampleVectors.clear();
    m_LabelVectors.clear();
    for (unsigned i = 0; i < (unsigned)X.cols(); ++i) {
        VectorType v(X.rows());
        for (unsigned k = 0; k < (unsigned)X.rows(); ++k) v[k] = X(k, i);
        m_SampleVectors.push_back(v);
    }
    for (unsigned i = 0; i < (unsigned)Y.cols(); ++i) {
        VectorType v(Y.rows());
        for (unsigned k = 0; k < (unsigned)Y.rows(); ++k) v[k] = Y(k, i);
        m_LabelVectors.push_back(v);
    }

    std::ifstream pf(pf_name.c_str());
    std::string line;
    if (std::getline(pf, line)) {
        std::stringstream ls(line);
        if (!(ls >> m_Sigma && ls >> m_InputDimension && ls >> m_OutputDimension && ls >> m_EfficientStorage &&
              ls >> debug))
            throw std::string("GaussianProcess::Load: parameter file is corrupt");
        std::string kernel_string;
        ls >> kernel_string;
        m_Kernel = KernelFactory<TScalarType>::GetKernel(kernel_string);
    }

    // install the loaded state on the device WITHOUT re-factorising (m_Initialized = true, reference :267)
    detail::DeviceModel& dev = Device();
    std::vector<int32_t> ops;
    std::vector<double> params;
    m_Kernel->Lower(ops, params);
    dev.Check(gprb200_set_kernel(dev.h, ops.data(), (int)ops.size(), params.data(), (int)params.size()));
    dev.sent_ops = ops;
    dev.sent_params = params;
    const size_t N = m_SampleVectors.size(), d = m_InputDimension, M = m_OutputDimension;
    std::vector<double> Xr(N * d), Yr(N * M);
    for (size_t i = 0; i < N; ++i) {
        for (size_t k = 0; k < d; ++k) Xr[i * d + k] = m_SampleVectors[i][k];
        for (size_t k = 0; k < M; ++k) Yr[i * M + k] = m_LabelVectors[i][k];
    }
    Staging<TScalarType> sa, sc;
    const double* pa = sa.In(m_RegressionVectors.data(), (size_t)m_RegressionVectors.size());
    const bool has_core = m_CoreMatrix.rows() > 0;
    const double* pc = has_core ? sc.In(m_CoreMatrix.data(), (size_t)m_CoreMatrix.size()) : nullptr;
    dev.Check(gprb200_set_state(dev.h, Xr.data(), (int64_t)N, (int)d, Yr.data(), pa, (int64_t)M, pc, (double)m_Sigma));
    dev.sent_sigma = (double)m_Sigma;
    m_TrainDirty = false;
    m_AlphaMirrored = true;
    m_CoreMirrored = true;
    m_CoreSize = (unsigned)m_CoreMatrix.diagonalSize();
    m_Initialized = true;
}

template <class TScalarType>
void GaussianProcess<TScalarType>::ToString() const {
    std::cout << "---------------------------------------" << std::endl;
    std::cout << "Gaussian Process" << std::endl;
    std::cout << " - initialized:\t\t" << m_Initialized << std::endl;
    std::cout << " - # samples:\t\t" << m_SampleVectors.size() << std::endl;
    std::cout << " - # labels:\t\t" << m_LabelVectors.size() << std::endl;
    std::cout << " - noise:\t\t" << m_Sigma << std::endl;
    std::cout << " - input dimension:\t" << m_InputDimension << std::endl;
    std::cout << " - output dimension:\t" << m_OutputDimension << std::endl;
    std::cout << std::endl;
    std::cout << " - Kernel:" << std::endl;
    std::cout << "       - Type:\t\t" << m_Kernel->ToString() << std::endl;
    std::cout << "       - Parameter:\t";
    for (const std::string& p : m_Kernel->GetStringParameters()) std::cout << p << ", ";
    std::cout << std::endl;
    std::cout << "---------------------------------------" << std::endl;
}

// reference :292-360: exact equality of alpha, samples, labels, kernel and flags; core matrix SIZE only
template <class TScalarType>
bool GaussianProcess<TScalarType>::operator==(const GaussianProcess<TScalarType>& b) const {
    struct Say {
        bool on;
        bool operator()(const char* why) const {
            if (on) std::cout << "GaussianProcess::comparison: " << why << std::endl;
            return false;
        }
    } differ = {debug};
    if (m_Initialized) MirrorRegressionVectors();
    if (b.m_Initialized) b.MirrorRegressionVectors();
    if (m_RegressionVectors.rows() != b.m_RegressionVectors.rows() ||
        m_RegressionVectors.cols() != b.m_RegressionVectors.cols())
        return differ("regression vectors not equal.");
    if (m_RegressionVectors.size() > 0 && (m_RegressionVectors - b.m_RegressionVectors).norm() > 0)
        return differ("regression vectors not equal.");
    if (m_CoreSize != b.m_CoreSize) return differ("core matrices not equal.");
    if (m_SampleVectors.size() != b.m_SampleVectors.size()) return differ("number of sample vectors not equal.");
    for (size_t i = 0; i < m_SampleVectors.size(); ++i)
        if ((m_SampleVectors[i] - b.m_SampleVectors[i]).norm() > 0) return differ("sample vectors not equal.");
    if (m_LabelVectors.size() != b.m_LabelVectors.size()) return differ("number of label vectors not equal.");
    for (size_t i = 0; i < m_LabelVectors.size(); ++i)
        if ((m_LabelVectors[i] - b.m_LabelVectors[i]).norm() > 0) return differ("label vectors not equal.");
    if (*m_Kernel.get() != *b.m_Kernel.get()) return differ("kernel not equal.");
    if (m_Sigma != b.m_Sigma) return differ("sigma not equal.");
    if (m_Initialized != b.m_Initialized) return differ("initialization state not equal.");
    if (m_InputDimension != b.m_InputDimension) return differ("input dimension not equal.");
    if (m_OutputDimension != b.m_OutputDimension) return differ("output dimension not equal.");
    if (m_EfficientStorage != b.m_EfficientStorage) return differ("efficient storage setting not equal.");
    if (debug != b.debug) return differ("debug state not equal.");
    if (debug) std::cout << "GaussianProcess::comparison: is equal!" << std::endl;
    return true;
}

// ---- protected: what Likelihood / subclasses reach -----------------------------------------------------
namespace {
// a throw-away device model for Gram-type computations on sample lists other than the GP's own
template <class T>
struct ScratchDevice {
    detail::DeviceModel dev;
    ScratchDevice(int device, const Kernel<T>& kernel, const std::vector<Eigen::Matrix<T, Eigen::Dynamic, 1> >& samples)
        : dev(device) {
        std::vector<int32_t> ops;
        std::vector<double> params;
        kernel.Lower(ops, params);
        dev.Check(gprb200_set_kernel(dev.h, ops.data(), (int)ops.size(), params.data(), (int)params.size()));
        const size_t N = samples.size(), d = N ? (size_t)samples[0].size() : 0;
        if (N == 0 || d == 0) throw std::string("GaussianProcess: no input samples defined");
        std::vector<double> X(N * d);
        for (size_t i = 0; i < N; ++i)
            for (size_t k = 0; k < d; ++k) X[i * d + k] = samples[i][k];
        dev.Check(gprb200_set_train(dev.h, X.data(), (int64_t)N, (int)d, nullptr, 0, 0.0));
    }
};
}  // namespace

// reference :363-372 / :384-402 (K without the noise term; throws if K is not finite)
template <class TScalarType>
void GaussianProcess<TScalarType>::ComputeKernelMatrix(MatrixType& M) const {
    if (debug) std::cout << "GaussianProcess::ComputeKernelMatrix: building kernel matrix... " << std::flush;
    SyncModelToDevice();
    detail::DeviceModel& dev = Device();
    const size_t N = m_SampleVectors.size();
    M.resize(N, N);
    Staging<TScalarType> st;
    double* p = st.Out(M.data(), N * N);
    dev.Check(gprb200_gram(dev.h, p, 0));
    st.Finish(M.data(), N * N);
    if (debug) std::cout << "[done]" << std::endl;
}

template <class TScalarType>
void GaussianProcess<TScalarType>::ComputeKernelMatrixInternal(MatrixType& M, const VectorListType& samples) const {
    ScratchDevice<TScalarType> s(m_DeviceIndex < 0 ? 0 : m_DeviceIndex, *m_Kernel, samples);
    const size_t N = samples.size();
    M.resize(N, N);
    Staging<TScalarType> st;
    double* p = st.Out(M.data(), N * N);
    s.dev.Check(gprb200_gram(s.dev.h, p, 0));
    st.Finish(M.data(), N * N);
}

// reference :375-381
template <class TScalarType>
void GaussianProcess<TScalarType>::AddNoiseToKernelMatrix(MatrixType& M) const {
    const TScalarType s2 = m_Sigma * m_Sigma;
    for (unsigned i = 0; i < (unsigned)M.rows(); ++i) M(i, i) += s2;
}

// reference :405-457 — O(N) host loops over the scalar kernel (used by the sparse likelihood only)
template <class TScalarType>
TScalarType GaussianProcess<TScalarType>::ComputeKernelMatrixTrace() const {
    return ComputeKernelMatrixTraceInternal(m_SampleVectors);
}
template <class TScalarType>
TScalarType GaussianProcess<TScalarType>::ComputeKernelMatrixTraceInternal(const VectorListType& samples) const {
    TScalarType t = 0;
    for (size_t i = 0; i < samples.size(); ++i) t += (*m_Kernel)(samples[i], samples[i]);
    return t;
}
template <class TScalarType>
typename GaussianProcess<TScalarType>::VectorType GaussianProcess<TScalarType>::ComputeDerivativeKernelMatrixTrace() const {
    return ComputeDerivativeKernelMatrixTraceInternal(m_SampleVectors);
}
template <class TScalarType>
typename GaussianProcess<TScalarType>::VectorType GaussianProcess<TScalarType>::ComputeDerivativeKernelMatrixTraceInternal(
    const VectorListType& samples) const {
    VectorType t = VectorType::Zero(m_Kernel->GetNumberOfParameters());
    for (size_t i = 0; i < samples.size(); ++i) t = t + m_Kernel->GetDerivative(samples[i], samples[i]);
    return t;
}

// reference :460-495 — the (P N) x N stack of dK/dtheta_p, block p at rows [pN, (p+1)N)
template <class TScalarType>
void GaussianProcess<TScalarType>::ComputeDerivativeKernelMatrix(MatrixType& M) const {
    SyncModelToDevice();
    detail::DeviceModel& dev = Device();
    const size_t N = m_SampleVectors.size(), P = m_Kernel->GetNumberOfParameters();
    M.resize(P * N, N);
    Staging<TScalarType> st;
    double* p = st.Out(M.data(), P * N * N);
    dev.Check(gprb200_dgram(dev.h, p));
    st.Finish(M.data(), P * N * N);
}

template <class TScalarType>
void GaussianProcess<TScalarType>::ComputeDerivativeKernelMatrixInternal(MatrixType& M, const VectorListType& samples) const {
    ScratchDevice<TScalarType> s(m_DeviceIndex < 0 ? 0 : m_DeviceIndex, *m_Kernel, samples);
    const size_t N = samples.size(), P = m_Kernel->GetNumberOfParameters();
    M.resize(P * N, N);
    Staging<TScalarType> st;
    double* p = st.Out(M.data(), P * N * N);
    s.dev.Check(gprb200_dgram(s.dev.h, p));
    st.Finish(M.data(), P * N * N);
}

// reference :498-510 — C = inv(K + sigma^2 I) with the CURRENT kernel parameters
template <class TScalarType>
void GaussianProcess<TScalarType>::ComputeCoreMatrix(MatrixType& C) const {
    ComputeCoreMatrixWithDeterminant(C);
}

// reference :513-528 — also det(K + sigma^2 I); here exp(log det) from the Cholesky diagonal in long double
template <class TScalarType>
typename GaussianProcess<TScalarType>::HighPrecisionType GaussianProcess<TScalarType>::ComputeCoreMatrixWithDeterminant(
    MatrixType& C) const {
    SyncModelToDevice();
    detail::DeviceModel& dev = Device();
    dev.Check(gprb200_factorize(dev.h));
    const size_t N = m_SampleVectors.size();
    C.resize(N, N);
    Staging<TScalarType> st;
    double* p = st.Out(C.data(), N * N);
    dev.Check(gprb200_get_core(dev.h, p));
    st.Finish(C.data(), N * N);
    double logdet = 0.0;
    dev.Check(gprb200_get_logdet(dev.h, &logdet));
    // the factorisation above replaced the one alpha was solved with
    m_AlphaMirrored = false;
    return std::exp(static_cast<HighPrecisionType>(logdet));
}

// reference :531-618 — every InversionMethod is the same Cholesky-based inverse here; K must be SPD
template <class TScalarType>
typename GaussianProcess<TScalarType>::MatrixType GaussianProcess<TScalarType>::InvertKernelMatrix(
    const MatrixType& K, InversionMethod, bool) const {
    if (K.rows() != K.cols()) throw std::string("GaussianProcess::InvertKernelMatrix: matrix is not square.");
    detail::DeviceModel& dev = Device();
    const size_t n = K.rows();
    MatrixType C(n, n);
    Staging<TScalarType> sk, sc;
    const double* pk = sk.In(K.data(), n * n);
    double* pc = sc.Out(C.data(), n * n);
    dev.Check(gprb200_spd_inverse(dev.h, pk, (int64_t)n, pc, nullptr));
    sc.Finish(C.data(), n * n);
    return C;
}

// reference :621-639
template <class TScalarType>
void GaussianProcess<TScalarType>::ComputeLabelMatrix(MatrixType& Y) const {
    ComputeLabelMatrixInternal(Y, m_LabelVectors);
}
template <class TScalarType>
void GaussianProcess<TScalarType>::ComputeLabelMatrixInternal(MatrixType& Y, const VectorListType& labels) const {
    if (labels.size() == 0) throw std::string("GaussianProcess::ComputeLabelMatrixInternal: no ouput labels defined.");
    const size_t n = labels.size(), m = labels[0].size();
    Y.resize(n, m);
    for (size_t i = 0; i < n; ++i)
        for (size_t k = 0; k < m; ++k) Y(i, k) = labels[i][k];
}

// reference :675-693
template <class TScalarType>
void GaussianProcess<TScalarType>::ComputeKernelVector(const VectorType& x, VectorType& Kx) const {
    if (!m_Initialized)
        throw std::string("GaussianProcess::ComputeKernelVectorInternal: gaussian process is not initialized.");
    SyncModelToDevice();
    detail::DeviceModel& dev = Device();
    const size_t N = m_SampleVectors.size();
    Kx.resize(N);
    Staging<TScalarType> sx, sk;
    const double* px = sx.In(x.data(), m_InputDimension);
    double* pk = sk.Out(Kx.data(), N);
    dev.Check(gprb200_kernel_vector(dev.h, px, 1, pk));
    sk.Finish(Kx.data(), N);
}
template <class TScalarType>
void GaussianProcess<TScalarType>::ComputeKernelVectorInternal(const VectorType& x, VectorType& Kx,
                                                               const VectorListType& samples) const {
    if (!m_Initialized)
        throw std::string("GaussianProcess::ComputeKernelVectorInternal: gaussian process is not initialized.");
    Kx.resize(samples.size());  // arbitrary sample list: N scalar kernel evaluations on the host
    for (size_t i = 0; i < samples.size(); ++i) Kx[i] = (*m_Kernel)(x, samples[i]);
}

// reference :697-706
template <class TScalarType>
void GaussianProcess<TScalarType>::ComputeDifferenceMatrix(const VectorType& x, MatrixType& X) const {
    const size_t n = m_SampleVectors.size(), d = x.size();
    X.resize(n, d);
    for (size_t i = 0; i < n; ++i)
        for (size_t k = 0; k < d; ++k) X(i, k) = x[k] - m_SampleVectors[i][k];
}

// reference :709-724
template <class TScalarType>
void GaussianProcess<TScalarType>::CheckInputDimension(const VectorType& x, std::string msg_prefix) const {
    if ((unsigned)x.size() != m_InputDimension) {
        std::stringstream msg;
        msg << msg_prefix << "dimension of input vector (" << x.size() << ") does not correspond to the input dimension ("
            << m_InputDimension << ").";
        throw std::string(msg.str());
    }
}
template <class TScalarType>
void GaussianProcess<TScalarType>::CheckOutputDimension(const VectorType& y, std::string msg_prefix) const {
    if ((unsigned)y.size() != m_OutputDimension) {
        std::stringstream msg;
        msg << msg_prefix << "dimension of output vector (" << y.size() << ") does not correspond to the output dimension ("
            << m_OutputDimension << ").";
        throw std::string(msg.str());
    }
}

// GaussianLogLikelihood on the device (include/Likelihood.h:166-344)
template <class TScalarType>
void GaussianProcess<TScalarType>::DeviceLogLikelihood(VectorType* value, VectorType* grad, MatrixType* jacobian) const {
    if (m_LabelVectors.size() == 0)
        throw std::string("GaussianProcess::ComputeLabelMatrixInternal: no ouput labels defined.");
    SyncModelToDevice();
    detail::DeviceModel& dev = Device();
    const size_t M = m_OutputDimension, P = m_Kernel->GetNumberOfParameters();
    std::vector<double> v(M), g(P), j(M * P);
    const int mode = (m_DeterminantMode == TrueLogDeterminant) ? GPRB200_LOGLIK_TRUE : GPRB200_LOGLIK_REFERENCE_CLAMP;
    dev.Check(gprb200_loglik(dev.h, mode, value ? v.data() : nullptr, grad ? g.data() : nullptr,
                             jacobian ? j.data() : nullptr));
    if (value) {
        value->resize(M);
        for (size_t i = 0; i < M; ++i) (*value)[i] = static_cast<TScalarType>(v[i]);
    }
    if (grad) {
        grad->resize(P);
        for (size_t i = 0; i < P; ++i) (*grad)[i] = static_cast<TScalarType>(g[i]);
    }
    if (jacobian) {
        jacobian->resize(M, P);
        for (size_t i = 0; i < M; ++i)
            for (size_t p = 0; p < P; ++p) (*jacobian)(i, p) = static_cast<TScalarType>(j[i * P + p]);
    }
    // the device now holds alpha for the CURRENT kernel parameters; the host mirror is stale
    m_AlphaMirrored = false;
}

template class GaussianProcess<float>;
template class GaussianProcess<double>;

}  // namespace gpr
