// gpr::GaussianProcess<T> — the reference's public class (include/GaussianProcess.h:33-329) over the
// B200 CUDA library. Same names, signatures, typedefs, protected virtuals and error convention
// (throw std::string), so apps/GaussianProcessLearn.cpp, apps/GaussianProcessPredict.cpp, the
// Likelihood classes and subclasses recompile unchanged; the work underneath goes through the
// C-ABI of include/gprb200.h (no CPU implementation of the path lives here):
//
//   Initialize -> ComputeRegressionVectors   reference lib/GaussianProcess.cpp:118-130, :642-672
//       Gram + sigma^2 I, fp64 Cholesky, alpha = (K + sigma^2 I)^-1 Y        gprb200_learn
//   Predict / PredictDerivative               :55-61, :65-81                  gprb200_predict[_derivative]
//   operator()(x,y) / GetCredibleInterval     :85-99, :103-114                gprb200_posterior_cov
//   Save / Load                               :134-268                        MatrixIO + gprb200_set_state
//
// State lives on the GPU. m_RegressionVectors / m_CoreMatrix are host MIRRORS: alpha is mirrored
// eagerly while it is small and on demand otherwise, the explicit inverse C = (K + sigma^2 I)^-1 is
// only formed when something observes it (Save without efficient storage, the likelihood gradient,
// ComputeCoreMatrix). All four InversionMethod values select the same Cholesky factorisation.
#ifndef GPRB200_CXX_GAUSSIAN_PROCESS_H
#define GPRB200_CXX_GAUSSIAN_PROCESS_H

#include <iostream>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include <Eigen/Dense>

#include "Kernel.h"

namespace gpr {

template <class TScalarType>
class Likelihood;

namespace detail {
struct DistModel;    // opaque: libgprb200_dist.so handle of this rank (src/GaussianProcess.cpp)
struct DeviceModel;  // opaque: gprb200 handle + what has been sent to it (src/GaussianProcess.cpp)
}

template <class TScalarType>
class GaussianProcess {
public:
    typedef GaussianProcess Self;
    typedef std::shared_ptr<Self> Pointer;
    typedef Kernel<TScalarType> KernelType;
    typedef typename KernelType::Pointer KernelTypePointer;

    typedef Eigen::Matrix<TScalarType, Eigen::Dynamic, 1> VectorType;
    typedef Eigen::Matrix<TScalarType, Eigen::Dynamic, Eigen::Dynamic, Eigen::RowMajor> MatrixType;
    typedef Eigen::DiagonalMatrix<TScalarType, Eigen::Dynamic> DiagMatrixType;

    typedef std::vector<VectorType> VectorListType;
    typedef std::vector<MatrixType> MatrixListType;

    typedef long double HighPrecisionType;

    // kept for source compatibility; every value runs the same fp64 Cholesky on the GPU
    typedef enum { FullPivotLU = 0, JacobiSVD = 1, BDCSVD = 2, SelfAdjointEigenSolver = 3 } InversionMethod;

    // how GaussianLogLikelihood forms its complexity term: the reference's clamped determinant
    // (include/Likelihood.h:180-188) or the true -0.5 log det from the Cholesky diagonal
    typedef enum { ReferenceClamp = 0, TrueLogDeterminant = 1 } DeterminantMode;

    void AddSample(const VectorType& x, const VectorType& y);
    virtual VectorType Predict(const VectorType& x);
    virtual VectorType PredictDerivative(const VectorType& x, MatrixType& D);
    virtual TScalarType operator()(const VectorType& x, const VectorType& y);
    TScalarType GetCredibleInterval(const VectorType& x);
    virtual void Initialize();

    // Additive batch entry point (not in the reference): T test points per call, one per row of Xs.
    // mean is T x M; ci (optional) receives GetCredibleInterval for every row.
    virtual void PredictBatch(const MatrixType& Xs, MatrixType& mean, VectorType* ci = 0);

    GaussianProcess(KernelTypePointer kernel);
    virtual ~GaussianProcess();

    const KernelTypePointer GetKernel() { return m_Kernel; }
    void SetKernel(KernelTypePointer k) {
        m_Kernel = k;
        m_Initialized = false;
    }

    void DebugOn() { debug = true; }
    virtual unsigned GetNumberOfSamples() const { return m_SampleVectors.size(); }
    TScalarType GetSigma() const { return m_Sigma; }
    TScalarType GetSigmaSquared() const { return m_Sigma * m_Sigma; }
    void SetSigma(TScalarType sigma) {
        m_Sigma = sigma;
        m_Initialized = false;
    }
    virtual unsigned GetNumberOfInputDimensions() const { return m_InputDimension; }

    virtual void SetInversionMethod(InversionMethod m) { m_InvMethod = m; }
    virtual InversionMethod GetInversionMethod() { return m_InvMethod; }

    bool GetEfficientStorage() { return m_EfficientStorage; }
    void SetEfficientStorage(bool s) { m_EfficientStorage = s; }

    void SetDeterminantMode(DeterminantMode m) { m_DeterminantMode = m; }
    DeterminantMode GetDeterminantMode() const { return m_DeterminantMode; }
    // CUDA device this model lives on (default: $GPRB200_DEVICE or 0); call before the first use
    void SetDevice(int device);
    // ONE model over several GPUs (not in the reference, which is a single process; include/gprb200_dist.h). Call on
    // every rank - one process or thread per GPU, each with its own GaussianProcess holding the same samples - before
    // Initialize(): ComputeRegressionVectors then factors K + sigma^2 I on the Pr x Pc grid of GPUs (block-cyclic
    // Cholesky, NCCL) and every rank keeps the regression vectors, so Predict runs locally on each. nccl_id: the
    // GetDistributedId() bytes of rank 0, carried to the other ranks by the application (MPI_Bcast, a file, ...);
    // world == 1 needs none. Pr = Pc = 0: the library's default grid. libgprb200_dist.so is opened on first use, so
    // programs that never call this do not need NCCL. The Cholesky factor stays on the grid: operator()(x, y) and
    // GetCredibleInterval(x) then solve there and are COLLECTIVE (every rank makes the same call and gets the value).
    void SetDistributed(int rank, int world, const void* nccl_id, int Pr = 0, int Pc = 0);
    static std::vector<unsigned char> GetDistributedId();
    // milliseconds of the last Gram / Cholesky / solve / inverse / predict phases (CUDA events)
    std::vector<double> GetPhaseMilliseconds() const;
    // The gprb200_handle (include/gprb200.h) this model lives on, initialised and in sync with the host state.
    // Lets device-side companions (PCA.h: fused Predict -> GetReconstruction) work on the same GPU and stream.
    void* GetDeviceHandle();

    virtual void Save(std::string prefix);
    virtual void Load(std::string prefix);
    virtual void ToString() const;

    void Lock() { gp_lock.lock(); }
    void UnLock() { gp_lock.unlock(); }

    virtual bool operator==(const GaussianProcess<TScalarType>& b) const;
    virtual bool operator!=(const GaussianProcess<TScalarType>& b) const { return !operator==(b); }

protected:
    virtual void ComputeKernelMatrix(MatrixType& M) const;
    virtual void AddNoiseToKernelMatrix(MatrixType& M) const;
    virtual TScalarType ComputeKernelMatrixTrace() const;
    virtual VectorType ComputeDerivativeKernelMatrixTrace() const;
    virtual void ComputeDerivativeKernelMatrix(MatrixType& M) const;
    virtual void ComputeDerivativeKernelMatrixInternal(MatrixType& M, const VectorListType& samples) const;
    void ComputeCoreMatrix(MatrixType& C) const;
    HighPrecisionType ComputeCoreMatrixWithDeterminant(MatrixType& C) const;
    virtual MatrixType InvertKernelMatrix(const MatrixType& K, InversionMethod inv_method = FullPivotLU,
                                          bool stable = false) const;
    virtual void ComputeLabelMatrix(MatrixType& Y) const;
    virtual void ComputeRegressionVectors();
    virtual void ComputeKernelVector(const VectorType& x, VectorType& Kx) const;
    void ComputeDifferenceMatrix(const VectorType& x, MatrixType& X) const;
    void CheckInputDimension(const VectorType& x, std::string msg_prefix) const;
    void CheckOutputDimension(const VectorType& y, std::string msg_prefix) const;

    KernelTypePointer m_Kernel;
    TScalarType m_Sigma;

    VectorListType m_SampleVectors;
    VectorListType m_LabelVectors;
    mutable MatrixType m_RegressionVectors;  // host mirror of alpha (see MirrorRegressionVectors)
    mutable MatrixType m_CoreMatrix;         // host mirror of C (see MirrorCoreMatrix)

    bool m_Initialized;
    unsigned m_InputDimension;
    unsigned m_OutputDimension;
    InversionMethod m_InvMethod;
    bool m_EfficientStorage;

    std::mutex gp_lock;
    bool debug;

    void ComputeKernelMatrixInternal(MatrixType& M, const VectorListType& samples) const;
    TScalarType ComputeKernelMatrixTraceInternal(const VectorListType& samples) const;
    VectorType ComputeDerivativeKernelMatrixTraceInternal(const VectorListType& samples) const;
    void ComputeLabelMatrixInternal(MatrixType& Y, const VectorListType& labels) const;
    void ComputeKernelVectorInternal(const VectorType& x, VectorType& Kx, const VectorListType& samples) const;

    // ---- device side -------------------------------------------------------------------------
    // Brings the host mirrors up to date with the device (no-ops when they already are).
    void MirrorRegressionVectors() const;
    void MirrorCoreMatrix() const;
    // GaussianLogLikelihood on the device (include/Likelihood.h:166-344): any of value (M),
    // grad (P), jacobian (M x P) may be null. Re-factorises with the current kernel parameters.
    void DeviceLogLikelihood(VectorType* value, VectorType* grad, MatrixType* jacobian) const;

    // The gprb200_handle with kernel, samples and sigma in sync with the host state, WITHOUT learning the dense model
    // (subclasses that factor something else on the same samples: SparseGaussianProcess.h).
    void* SyncedDeviceHandle() const;

    DeterminantMode m_DeterminantMode;
    mutable unsigned m_CoreSize;  // what m_CoreMatrix.diagonalSize() is in the reference's state machine

private:
    GaussianProcess(const Self&);
    void operator=(const Self&);

    detail::DeviceModel& Device() const;
    void SyncModelToDevice() const;  // kernel program, samples/labels, sigma
    mutable std::unique_ptr<detail::DeviceModel> m_Device;
    std::shared_ptr<detail::DistModel> m_Dist;   // set by SetDistributed
    void ComputeRegressionVectorsDistributed();
    mutable bool m_TrainDirty;
    mutable bool m_AlphaMirrored;
    mutable bool m_CoreMirrored;
    int m_DeviceIndex;

    friend class Likelihood<TScalarType>;
};

}  // namespace gpr

#include "Likelihood.h"

#endif
