// PCA<T> — the reference's class (include/PCA.h:21-145) over the B200 library: same constructors, method names and
// argument meaning, so include/DataParser.h (ComputeFeaturesForTraining, inversePca, GetComputationTime) compiles
// against it unchanged. Data matrices are D x n with one sample per column. The basis lives on the GPU
// (gpr_b200/csrc/pca.cu through include/gprb200.h); no CPU implementation of the decomposition or of the
// projections lives here.
//
//   PCA(X, nFeatures)            mean, thin SVD of the mean-free data, sigma = s / sqrt(n)     gprb200_pca_fit
//   PCA(path, nFeatures)         Mean.bin / Sigma.bin / U.bin (MatrixIO)                       gprb200_pca_set
//   DimensionalityReduction[Fast] basis^T (X - mean)                                            gprb200_pca_project
//   GetReconstruction            basisInverse[:, :nFeatures] W + mean                          gprb200_pca_reconstruct
//
// Additive (not in the reference): a constructor argument that places the PCA on the GPU of an existing
// GaussianProcess, and PredictReconstruction(gp, Xs): gp->Predict for every row of Xs followed by GetReconstruction
// without the weights leaving the device (the per-frame path of apps/GaussianProcessPredict.cpp:188-203).
// Singular vectors are defined up to sign, as with Eigen's BDCSVD.
#pragma once

#ifndef PROJECT_PCA_H
#define PROJECT_PCA_H

#include <cmath>
#include <iostream>
#include <memory>
#include <string>
#include <vector>

#include <Eigen/Dense>

#include "GaussianProcess.h"
#include "MatrixIO.h"
#include "gprb200.h"

template <class TScalarType>
class PCA {
public:
    typedef gpr::GaussianProcess<TScalarType> GaussianProcessType;
    typedef std::shared_ptr<GaussianProcessType> GaussianProcessTypePointer;
    typedef typename GaussianProcessType::VectorType VectorType;
    typedef typename GaussianProcessType::MatrixType MatrixType;
    typedef typename std::shared_ptr<VectorType> VectorTypePointer;
    typedef typename std::shared_ptr<MatrixType> MatrixTypePointer;

    PCA(MatrixType& X, int nFeatures, GaussianProcessTypePointer gp = GaussianProcessTypePointer()) {
        Open(gp);
        m_nFeatures = nFeatures;
        std::vector<double> x = ToDouble(X.data(), (size_t)X.rows() * X.cols());
        Check(gprb200_pca_fit(m_pca, x.data(), X.rows(), X.cols(), nFeatures));
        m_D = X.rows();
        m_n = X.cols();
    }

    PCA(std::string path, int nFeatures, GaussianProcessTypePointer gp = GaussianProcessTypePointer()) {
        Open(gp);
        m_nFeatures = nFeatures;
        // Read PCA matrices from files (computed based on trainings data), reference include/PCA.h:54-61
        MatrixType mean = gpr::ReadMatrix<MatrixType>(path + "Mean.bin");
        MatrixType sigma = gpr::ReadMatrix<MatrixType>(path + "Sigma.bin");
        MatrixType U = gpr::ReadMatrix<MatrixType>(path + "U.bin");
        std::vector<double> m = ToDouble(mean.data(), (size_t)mean.size());
        std::vector<double> s = ToDouble(sigma.data(), (size_t)sigma.size());
        std::vector<double> u = ToDouble(U.data(), (size_t)U.size());
        if ((long)mean.size() != (long)U.rows() || (long)sigma.size() != (long)U.cols())
            throw std::string("PCA: Mean/Sigma/U files do not belong together");
        Check(gprb200_pca_set(m_pca, m.data(), s.data(), u.data(), U.rows(), U.cols(), nFeatures));
        m_D = U.rows();
        m_n = U.cols();
    }

    ~PCA() {
        if (m_pca) gprb200_pca_destroy(m_pca);
        if (m_ownHandle && m_handle) gprb200_destroy(m_handle);
    }

    // the compact transposed basis is built on the device when the basis is installed
    void PrecomputeTranspose() {}

    VectorType GetMean() {
        std::vector<double> m((size_t)m_D);
        Check(gprb200_pca_get(m_pca, m.data(), nullptr, nullptr));
        VectorType v(m_D);
        for (long i = 0; i < m_D; ++i) v[i] = (TScalarType)m[(size_t)i];
        return v;
    }

    VectorType GetEigenvalues() {
        std::vector<double> s((size_t)m_n);
        Check(gprb200_pca_get(m_pca, nullptr, s.data(), nullptr));
        VectorType v(m_n);
        for (long i = 0; i < m_n; ++i) v[i] = (TScalarType)s[(size_t)i];
        return v;
    }

    MatrixType GetMatrixU() {
        std::vector<double> u((size_t)m_D * m_n);
        Check(gprb200_pca_get(m_pca, nullptr, nullptr, u.data()));
        MatrixType U(m_D, m_n);
        for (size_t i = 0; i < u.size(); ++i) U.data()[i] = (TScalarType)u[i];
        return U;
    }

    // U diag(1/sigma), include/PCA.h:46, :83-90
    MatrixType GetBasis(int nFeatures = 0) {
        MatrixType U = GetMatrixU();
        VectorType s = GetEigenvalues();
        const long cols = (nFeatures > 0 && nFeatures < m_n) ? nFeatures : m_n;
        MatrixType B(m_D, cols);
        for (long i = 0; i < m_D; ++i)
            for (long k = 0; k < cols; ++k) B(i, k) = U(i, k) / s[k];
        return B;
    }

    // every mode: the reference's `nFeatures << features.rows()` (include/PCA.h:95) is true for any nFeatures > 0
    // with a non-zero shift result, so topRows(nFeatures) is what it returns when nFeatures > 0
    MatrixType DimensionalityReduction(MatrixType& X, int nFeatures = 0) {
        const long rows = (nFeatures > 0 && nFeatures <= m_n) ? nFeatures : m_n;
        return Project(X, rows);
    }

    MatrixType DimensionalityReductionFast(MatrixType& X) { return Project(X, m_nFeatures); }

    MatrixType GetReconstruction(MatrixType& weights) {
        std::vector<double> w = ToDouble(weights.data(), (size_t)weights.size());
        std::vector<double> out((size_t)m_D * weights.cols());
        Check(gprb200_pca_reconstruct(m_pca, w.data(), weights.rows(), weights.cols(), out.data()));
        MatrixType X(m_D, weights.cols());
        for (size_t i = 0; i < out.size(); ++i) X.data()[i] = (TScalarType)out[i];
        return X;
    }

    // cumulative sigma, normalised: the intent of include/PCA.h:117-124 (the reference indexes the vector m_sigma
    // as (i, i), which is out of bounds for i >= 1)
    VectorType GetExplainedVariance() {
        VectorType s = GetEigenvalues();
        VectorType c = VectorType::Zero(m_n);
        TScalarType acc = 0;
        for (long i = 0; i < m_n; ++i) {
            acc += s[i];
            c[i] = acc;
        }
        for (long i = 0; i < m_n; ++i) c[i] /= acc;
        return c;
    }

    void WriteMatricesToFile(std::string path) {
        MatrixType mean = GetMean(), sigma = GetEigenvalues(), U = GetMatrixU();
        gpr::WriteMatrix<MatrixType>(mean, path + "Mean.bin");
        gpr::WriteMatrix<MatrixType>(sigma, path + "Sigma.bin");
        gpr::WriteMatrix<MatrixType>(U, path + "U.bin");
    }

    // ---- additive ------------------------------------------------------------------------------------------
    // fields (D x T) for the T test points in the rows of Xs, through `gp` (which must be the GaussianProcess this
    // PCA was constructed with and must predict nFeatures outputs). weights (optional): T x nFeatures.
    MatrixType PredictReconstruction(GaussianProcessTypePointer gp, const MatrixType& Xs, MatrixType* weights = 0) {
        if (!gp || (gprb200_handle)gp->GetDeviceHandle() != m_handle)
            throw std::string("PCA::PredictReconstruction: construct the PCA with this GaussianProcess");
        std::vector<double> xs = ToDouble(Xs.data(), (size_t)Xs.size());
        std::vector<double> out((size_t)m_D * Xs.rows()), w;
        if (weights) w.resize((size_t)Xs.rows() * m_nFeatures);
        Check(gprb200_predict_reconstruct(m_handle, m_pca, xs.data(), Xs.rows(), out.data(), weights ? w.data() : nullptr));
        MatrixType X(m_D, Xs.rows());
        for (size_t i = 0; i < out.size(); ++i) X.data()[i] = (TScalarType)out[i];
        if (weights) {
            weights->resize(Xs.rows(), m_nFeatures);
            for (size_t i = 0; i < w.size(); ++i) weights->data()[i] = (TScalarType)w[i];
        }
        return X;
    }

    double GetFitMilliseconds() {
        double v = 0;
        gprb200_pca_get_stat(m_pca, "fit_ms", &v);
        return v;
    }

private:
    PCA(const PCA&);
    void operator=(const PCA&);

    void Open(GaussianProcessTypePointer gp) {
        m_pca = nullptr;
        m_handle = nullptr;
        m_ownHandle = false;
        if (gp) {
            m_handle = (gprb200_handle)gp->GetDeviceHandle();
        } else {
            const char* env = std::getenv("GPRB200_DEVICE");
            if (gprb200_create(&m_handle, env ? std::atoi(env) : 0) != 0) throw std::string("PCA: ") + gprb200_last_error(nullptr);
            m_ownHandle = true;
        }
        if (gprb200_pca_create(m_handle, &m_pca) != 0) throw std::string(gprb200_last_error(m_handle));
    }
    void Check(int rc) const {
        if (rc != 0) throw std::string(gprb200_last_error(m_handle));
    }
    static std::vector<double> ToDouble(const TScalarType* p, size_t count) {
        std::vector<double> v(count);
        for (size_t i = 0; i < count; ++i) v[i] = (double)p[i];
        return v;
    }
    MatrixType Project(MatrixType& X, long rows) {
        if ((long)X.rows() != m_D) throw std::string("PCA::DimensionalityReduction: dimension mismatch");
        std::vector<double> x = ToDouble(X.data(), (size_t)X.size());
        std::vector<double> f((size_t)rows * X.cols());
        Check(gprb200_pca_project(m_pca, x.data(), X.cols(), rows, f.data()));
        MatrixType F(rows, X.cols());
        for (size_t i = 0; i < f.size(); ++i) F.data()[i] = (TScalarType)f[i];
        return F;
    }

    int m_nFeatures;
    long m_D, m_n;
    gprb200_handle m_handle;
    gprb200_pca_handle m_pca;
    bool m_ownHandle;
};

#endif  // PROJECT_PCA_H
