// Host <-> device movement of the label / regression matrices (G1, G5, G15 of SURVEY.md 8(a)).
//
// The reference keeps samples and labels as std::vector<VectorType> (lib/GaussianProcess.cpp:48-49) and gathers them
// into one N x M matrix (ComputeLabelMatrixInternal, :621-639); at the motion-model size (N = 4096, M = 786432) that
// matrix is 25.8 GB and moving it dominates everything but the solve. Here rows go from wherever they live (one
// pointer per row, or one contiguous block) through a ring of pinned staging buffers: worker threads copy piece k+1
// into pinned memory while piece k crosses PCIe with cudaMemcpy2DAsync, so nothing pageable is ever handed to the
// driver, no second N x M host copy exists, and the transfer runs at the pinned rate. The way back (get_alpha for
// Save / the host mirror) is the same pipeline in the other direction.
#include "common.cuh"
#include <string.h>
#include <thread>
#include <vector>

namespace gprb200 {

namespace {

constexpr size_t PIECE_BYTES = 64u << 20;   // one staging buffer
constexpr int NTHREADS = 8;                 // copy workers per piece

int ensure_pinned(Ctx* c, int ring) {
    for (int k = 0; k < Ctx::NPIN; ++k) {
        if (!c->pin[ring][k]) GP_CUDA(c, cudaHostAlloc(&c->pin[ring][k], PIECE_BYTES, cudaHostAllocDefault));
        if (!c->pin_ev[ring][k]) GP_CUDA(c, cudaEventCreateWithFlags(&c->pin_ev[ring][k], cudaEventDisableTiming));
    }
    return 0;
}

// run f(t) for t in [0, n) on n threads (n small); the calling thread takes t = 0
template <class F>
void parallel(int n, F f) {
    std::vector<std::thread> th;
    for (int t = 1; t < n; ++t) th.emplace_back(f, t);
    f(0);
    for (auto& x : th) x.join();
}

struct Piece {
    int64_t r0, nr, c0, nc;
};

std::vector<Piece> pieces(int64_t rows, int64_t cols) {
    const int64_t cap = (int64_t)(PIECE_BYTES / sizeof(double));
    std::vector<Piece> v;
    const int64_t cw = cols < cap ? cols : cap;               // columns per piece
    const int64_t rh = cw < cols ? 1 : (cap / cw > 0 ? cap / cw : 1);   // rows per piece (1 when a row is split)
    for (int64_t r0 = 0; r0 < rows; r0 += rh)
        for (int64_t c0 = 0; c0 < cols; c0 += cw)
            v.push_back({r0, (rows - r0 < rh) ? rows - r0 : rh, c0, (cols - c0 < cw) ? cols - c0 : cw});
    return v;
}

// copy piece p between the staging buffer (pitch p.nc) and host rows; to_pinned selects the direction
void copy_piece(const Piece& p, double* buf, const double* const* rowptr, const double* flat, int64_t ldf, bool to_pinned) {
    const int nt = (p.nr * p.nc * (int64_t)sizeof(double) < (4 << 20)) ? 1 : NTHREADS;
    parallel(nt, [&](int t) {
        if (p.nr >= nt) {   // split by rows
            for (int64_t r = t; r < p.nr; r += nt) {
                double* h = rowptr ? const_cast<double*>(rowptr[p.r0 + r]) + p.c0
                                   : const_cast<double*>(flat) + (p.r0 + r) * ldf + p.c0;
                double* b = buf + r * p.nc;
                if (to_pinned) memcpy(b, h, sizeof(double) * p.nc);
                else memcpy(h, b, sizeof(double) * p.nc);
            }
        } else {            // few long rows: split every row by columns
            const int64_t seg = (p.nc + nt - 1) / nt, a = t * seg, e = (a + seg < p.nc) ? a + seg : p.nc;
            if (a >= e) return;
            for (int64_t r = 0; r < p.nr; ++r) {
                double* h = rowptr ? const_cast<double*>(rowptr[p.r0 + r]) + p.c0
                                   : const_cast<double*>(flat) + (p.r0 + r) * ldf + p.c0;
                double* b = buf + r * p.nc;
                if (to_pinned) memcpy(b + a, h + a, sizeof(double) * (e - a));
                else memcpy(h + a, b + a, sizeof(double) * (e - a));
            }
        }
    });
}

}  // namespace

// dst (device, rows x cols, leading dimension ldd) <- host rows; rowptr (one pointer per row) or flat (leading
// dimension ldf), each read from column hoff. Asynchronous on st except for the staging-buffer reuse waits; the
// caller synchronises. Ring 0.
int stream_h2d(Ctx* c, double* dst, int64_t ldd, const double* const* rowptr, const double* flat, int64_t ldf, int64_t rows,
               int64_t cols, cudaStream_t st, int64_t hoff) {
    if (rows <= 0 || cols <= 0) return 0;
    if (!st) st = c->stream;
    GP_TRY(ensure_pinned(c, 0));
    // a previous call may still be reading the ring (its copies are asynchronous)
    for (int k = 0; k < Ctx::NPIN; ++k) GP_CUDA(c, cudaEventSynchronize(c->pin_ev[0][k]));
    const std::vector<Piece> ps = pieces(rows, cols);
    for (size_t k = 0; k < ps.size(); ++k) {
        const int b = (int)(k % Ctx::NPIN);
        if (k >= (size_t)Ctx::NPIN) GP_CUDA(c, cudaEventSynchronize(c->pin_ev[0][b]));   // its previous upload has left the buffer
        Piece p = ps[k];
        p.c0 += hoff;
        copy_piece(p, (double*)c->pin[0][b], rowptr, flat, ldf, true);
        GP_CUDA(c, cudaMemcpy2DAsync(dst + p.r0 * ldd + ps[k].c0, sizeof(double) * ldd, c->pin[0][b], sizeof(double) * p.nc,
                                     sizeof(double) * p.nc, (size_t)p.nr, cudaMemcpyHostToDevice, st));
        GP_CUDA(c, cudaEventRecord(c->pin_ev[0][b], st));
    }
    return 0;
}

// host (flat, leading dimension ldf) <- src (device, rows x cols, leading dimension lds); returns when the host has it
int stream_d2h(Ctx* c, double* flat, int64_t ldf, const double* src, int64_t lds, int64_t rows, int64_t cols,
               cudaStream_t st) {
    if (rows <= 0 || cols <= 0) return 0;
    if (!st) st = c->stream;
    GP_TRY(ensure_pinned(c, 1));
    const std::vector<Piece> ps = pieces(rows, cols);
    auto issue = [&](size_t k) -> int {
        const int b = (int)(k % Ctx::NPIN);
        const Piece& p = ps[k];
        GP_CUDA(c, cudaMemcpy2DAsync(c->pin[1][b], sizeof(double) * p.nc, src + p.r0 * lds + p.c0, sizeof(double) * lds,
                                     sizeof(double) * p.nc, (size_t)p.nr, cudaMemcpyDeviceToHost, st));
        GP_CUDA(c, cudaEventRecord(c->pin_ev[1][b], st));
        return 0;
    };
    const size_t ahead = Ctx::NPIN - 1;   // downloads in flight while the host empties a buffer
    for (size_t k = 0; k < ps.size() && k < ahead; ++k) GP_TRY(issue(k));
    for (size_t k = 0; k < ps.size(); ++k) {
        if (k + ahead < ps.size()) GP_TRY(issue(k + ahead));
        const int b = (int)(k % Ctx::NPIN);
        GP_CUDA(c, cudaEventSynchronize(c->pin_ev[1][b]));
        copy_piece(ps[k], (double*)c->pin[1][b], nullptr, flat, ldf, false);
    }
    return 0;
}

void hostio_destroy(Ctx* c) {
    for (int r = 0; r < 2; ++r) {
        for (int k = 0; k < Ctx::NPIN; ++k) {
            if (c->pin[r][k]) cudaFreeHost(c->pin[r][k]);
            if (c->pin_ev[r][k]) cudaEventDestroy(c->pin_ev[r][k]);
            c->pin[r][k] = nullptr;
            c->pin_ev[r][k] = nullptr;
        }
        if (c->io_stream[r]) cudaStreamDestroy(c->io_stream[r]);
        c->io_stream[r] = nullptr;
    }
}

}  // namespace gprb200
