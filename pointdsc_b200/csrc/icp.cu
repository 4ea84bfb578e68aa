// nvcc-flags: -fmad=false
// Point-to-point ICP on the key points: what evaluation/benchmark_utils.py:40-56 (icp_refine) does through
// open3d's registration_icp(source, target, 0.1, init, TransformationEstimationPointToPoint()) with the default
// ICPConvergenceCriteria (relative_fitness 1e-6, relative_rmse 1e-6, max_iteration 30), in float64 as open3d.
//
//   icp_grid_kernel    one CTA per pair: uniform grid over the target points (cell = max distance), histogram and scan in
//                      shared memory, points scattered into cell order
//   icp_assoc_kernel   one thread per source point: transform, nearest target point closer than the max distance
//                      (27 cells, ties to the lower index), per-CTA partial sums for fitness / RMSE / Umeyama
//   icp_update_kernel  one warp per pair: ordered reduction of the partials, convergence test against the previous
//                      association, rigid update (Kabsch, no scaling), T <- update * T
// The host enqueues (max_iteration + 1) x (assoc, update); pairs that have converged skip their work on the device, so
// nothing is read back between iterations.
#include <float.h>
#include <math.h>

#include "common.cuh"

namespace pdsc {

namespace {

constexpr int kIcpCells = 32768;     // cells per pair (128 KB of shared memory for the histogram)
constexpr int kGridThreads = 1024;
constexpr int kAssocThreads = 256;
constexpr int kSums = 17;            // n, sum d2, sum p (3), sum q (3), sum p q^T (9)

struct IcpGrid {
    double origin[3];
    double h;
    int dims[3];
    int ncells;
};

struct IcpState {
    double T[16];
    double prev_fitness, prev_rmse;
    double fitness, rmse;
    int done, updates, assocs, pad;
};

__device__ __forceinline__ double blk_minmax(double v, bool is_min, double* sh) {
    for (int o = 16; o > 0; o >>= 1) {
        double w = __shfl_xor_sync(0xffffffffu, v, o);
        v = is_min ? fmin(v, w) : fmax(v, w);
    }
    __syncthreads();
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x < 32) {
        v = sh[threadIdx.x];
        for (int o = 16; o > 0; o >>= 1) {
            double w = __shfl_xor_sync(0xffffffffu, v, o);
            v = is_min ? fmin(v, w) : fmax(v, w);
        }
        if (threadIdx.x == 0) sh[0] = v;
    }
    __syncthreads();
    return sh[0];
}

__device__ __forceinline__ int grid_cell(const IcpGrid& g, double x, double y, double z) {
    int cx = min(max((int)floor((x - g.origin[0]) / g.h), 0), g.dims[0] - 1);
    int cy = min(max((int)floor((y - g.origin[1]) / g.h), 0), g.dims[1] - 1);
    int cz = min(max((int)floor((z - g.origin[2]) / g.h), 0), g.dims[2] - 1);
    return (cz * g.dims[1] + cy) * g.dims[0] + cx;
}

__global__ void __launch_bounds__(kGridThreads) icp_grid_kernel(const float* __restrict__ tgt, const float* __restrict__ trans, int N,
                                                                double max_dist, IcpGrid* __restrict__ grids, int* __restrict__ cell_end,
                                                                double* __restrict__ sorted_xyz, int* __restrict__ sorted_idx,
                                                                IcpState* __restrict__ state) {
    extern __shared__ int cnt[];          // [kIcpCells]
    __shared__ double sh[32];
    __shared__ IcpGrid g;
    __shared__ int warp_tot[32];
    const int b = blockIdx.x, t = threadIdx.x;
    const float* P = tgt + (size_t)b * N * 3;
    double lo[3] = {DBL_MAX, DBL_MAX, DBL_MAX}, hi[3] = {-DBL_MAX, -DBL_MAX, -DBL_MAX};
    for (int i = t; i < N; i += kGridThreads)
        for (int c = 0; c < 3; ++c) {
            double v = (double)P[(size_t)i * 3 + c];
            lo[c] = fmin(lo[c], v);
            hi[c] = fmax(hi[c], v);
        }
    for (int c = 0; c < 3; ++c) {
        lo[c] = blk_minmax(lo[c], true, sh);
        hi[c] = blk_minmax(hi[c], false, sh);
    }
    if (t == 0) {
        g.h = max_dist;
        for (int c = 0; c < 3; ++c) g.origin[c] = lo[c];
        for (int it = 0; it < 400; ++it) {
            double cells = 1.0;
            for (int c = 0; c < 3; ++c) {
                double d = fmin(fmax(floor((hi[c] - lo[c]) / g.h) + 1.0, 1.0), 1.0e9);
                g.dims[c] = (int)d;
                cells *= d;
            }
            g.ncells = cells <= (double)kIcpCells ? (int)cells : 0;
            if (g.ncells) break;
            g.h *= 1.25;
        }
        grids[b] = g;
        IcpState s;
        for (int i = 0; i < 16; ++i) s.T[i] = (double)trans[(size_t)b * 16 + i];
        s.prev_fitness = s.prev_rmse = s.fitness = s.rmse = 0.0;
        s.done = 0; s.updates = 0; s.assocs = 0; s.pad = 0;
        state[b] = s;
    }
    __syncthreads();
    const int ncells = g.ncells;
    for (int c = t; c < ncells; c += kGridThreads) cnt[c] = 0;
    __syncthreads();
    for (int i = t; i < N; i += kGridThreads)
        atomicAdd(&cnt[grid_cell(g, (double)P[(size_t)i * 3], (double)P[(size_t)i * 3 + 1], (double)P[(size_t)i * 3 + 2])], 1);
    __syncthreads();
    // exclusive scan: each thread owns a contiguous chunk of cells
    const int chunk = (ncells + kGridThreads - 1) / kGridThreads;
    const int c0 = min(t * chunk, ncells), c1 = min(c0 + chunk, ncells);
    int local = 0;
    for (int c = c0; c < c1; ++c) local += cnt[c];
    int incl = local;
    const int lane = t & 31, warp = t >> 5;
    for (int o = 1; o < 32; o <<= 1) {
        int v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
    }
    if (lane == 31) warp_tot[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        int w = warp_tot[lane];
        for (int o = 1; o < 32; o <<= 1) {
            int v = __shfl_up_sync(0xffffffffu, w, o);
            if (lane >= o) w += v;
        }
        warp_tot[lane] = w;
    }
    __syncthreads();
    int run = incl - local + (warp ? warp_tot[warp - 1] : 0);
    for (int c = c0; c < c1; ++c) {
        int v = cnt[c];
        cnt[c] = run;          // start offset; the scatter's atomicAdd turns it into the end offset
        run += v;
    }
    __syncthreads();
    double* SX = sorted_xyz + (size_t)b * N * 3;
    int* SI = sorted_idx + (size_t)b * N;
    for (int i = t; i < N; i += kGridThreads) {
        const double x = (double)P[(size_t)i * 3], y = (double)P[(size_t)i * 3 + 1], z = (double)P[(size_t)i * 3 + 2];
        const int pos = atomicAdd(&cnt[grid_cell(g, x, y, z)], 1);
        SX[(size_t)pos * 3] = x; SX[(size_t)pos * 3 + 1] = y; SX[(size_t)pos * 3 + 2] = z;
        SI[pos] = i;
    }
    __syncthreads();
    for (int c = t; c < ncells; c += kGridThreads) cell_end[(size_t)b * kIcpCells + c] = cnt[c];
}

__global__ void __launch_bounds__(kAssocThreads) icp_assoc_kernel(const float* __restrict__ src, int N, double max_dist,
                                                                  const IcpGrid* __restrict__ grids, const int* __restrict__ cell_end,
                                                                  const double* __restrict__ sorted_xyz, const int* __restrict__ sorted_idx,
                                                                  const IcpState* __restrict__ state, double* __restrict__ partials,
                                                                  int32_t* __restrict__ corr_out) {
    const int b = blockIdx.y;
    if (state[b].done) return;
    __shared__ double red[kAssocThreads / 32][kSums];
    const int i = blockIdx.x * kAssocThreads + threadIdx.x;
    double v[kSums];
#pragma unroll
    for (int k = 0; k < kSums; ++k) v[k] = 0.0;
    if (i < N) {
        const double* T = state[b].T;
        const float* s = src + ((size_t)b * N + i) * 3;
        const double x = (double)s[0], y = (double)s[1], z = (double)s[2];
        const double px = T[0] * x + T[1] * y + T[2] * z + T[3];
        const double py = T[4] * x + T[5] * y + T[6] * z + T[7];
        const double pz = T[8] * x + T[9] * y + T[10] * z + T[11];
        const IcpGrid g = grids[b];
        const int cx = (int)floor((px - g.origin[0]) / g.h), cy = (int)floor((py - g.origin[1]) / g.h),
                  cz = (int)floor((pz - g.origin[2]) / g.h);
        const int* CE = cell_end + (size_t)b * kIcpCells;
        const double* SX = sorted_xyz + (size_t)b * N * 3;
        const int* SI = sorted_idx + (size_t)b * N;
        double best = max_dist * max_dist;
        int bj = -1;
        double qx = 0, qy = 0, qz = 0;
        const int x0 = max(cx - 1, 0), x1 = min(cx + 1, g.dims[0] - 1);
        if (x0 <= x1)
            for (int zz = max(cz - 1, 0); zz <= min(cz + 1, g.dims[2] - 1); ++zz)
                for (int yy = max(cy - 1, 0); yy <= min(cy + 1, g.dims[1] - 1); ++yy) {
                    const int row = (zz * g.dims[1] + yy) * g.dims[0];
                    const int cb = (row + x0) > 0 ? CE[row + x0 - 1] : 0, ce = CE[row + x1];
                    for (int c = cb; c < ce; ++c) {
                        const double ex = SX[(size_t)c * 3] - px, ey = SX[(size_t)c * 3 + 1] - py, ez = SX[(size_t)c * 3 + 2] - pz;
                        const double d2 = ex * ex + ey * ey + ez * ez;
                        if (d2 > best || !(d2 < max_dist * max_dist)) continue;
                        const int j = SI[c];
                        if (d2 < best || j < bj || bj < 0) {
                            best = d2; bj = j;
                            qx = SX[(size_t)c * 3]; qy = SX[(size_t)c * 3 + 1]; qz = SX[(size_t)c * 3 + 2];
                        }
                    }
                }
        if (corr_out) corr_out[(size_t)b * N + i] = bj;
        if (bj >= 0) {
            v[0] = 1.0; v[1] = best;
            v[2] = px; v[3] = py; v[4] = pz;
            v[5] = qx; v[6] = qy; v[7] = qz;
            v[8] = px * qx; v[9] = px * qy; v[10] = px * qz;
            v[11] = py * qx; v[12] = py * qy; v[13] = py * qz;
            v[14] = pz * qx; v[15] = pz * qy; v[16] = pz * qz;
        }
    }
    // fixed-shape reduction: the result does not depend on scheduling
#pragma unroll
    for (int k = 0; k < kSums; ++k) {
        double w = v[k];
        for (int o = 16; o > 0; o >>= 1) w += __shfl_xor_sync(0xffffffffu, w, o);
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5][k] = w;
    }
    __syncthreads();
    if (threadIdx.x < kSums) {
        double w = 0.0;
        for (int q = 0; q < kAssocThreads / 32; ++q) w += red[q][threadIdx.x];
        partials[((size_t)b * gridDim.x + blockIdx.x) * kSums + threadIdx.x] = w;
    }
}

__global__ void __launch_bounds__(32) icp_update_kernel(int N, int nblk, int max_iter, double rel_fitness, double rel_rmse,
                                                        const double* __restrict__ partials, IcpState* __restrict__ state,
                                                        float* __restrict__ trans_out, double* __restrict__ stats_out,
                                                        int32_t* __restrict__ iters_out) {
    const int b = blockIdx.x, lane = threadIdx.x;
    IcpState* S = state + b;
    if (S->done) return;
    double sum[kSums];
    for (int k = 0; k < kSums; ++k) {
        double w = 0.0;
        for (int q = lane; q < nblk; q += 32) w += partials[((size_t)b * nblk + q) * kSums + k];
        for (int o = 16; o > 0; o >>= 1) w += __shfl_xor_sync(0xffffffffu, w, o);
        sum[k] = w;
    }
    if (lane != 0) return;
    const double n = sum[0];
    const double fitness = n / (double)N;
    const double rmse = n > 0.0 ? sqrt(sum[1] / n) : 0.0;
    const int k = S->assocs;             // index of the association just reduced
    bool done = false;
    if (k > 0 && fabs(S->prev_fitness - fitness) < rel_fitness && fabs(S->prev_rmse - rmse) < rel_rmse) done = true;
    if (k >= max_iter) done = true;
    S->fitness = fitness; S->rmse = rmse;
    S->prev_fitness = fitness; S->prev_rmse = rmse;
    S->assocs = k + 1;
    if (!done) {
        // Umeyama without scaling: R = U diag(1,1,det) V^T of sum (q - mq)(p - mp)^T, t = mq - R mp
        double R[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}, tr[3] = {0, 0, 0};
        if (n > 0.0) {
            const double mp[3] = {sum[2] / n, sum[3] / n, sum[4] / n}, mq[3] = {sum[5] / n, sum[6] / n, sum[7] / n};
            double H[9];
            for (int r = 0; r < 3; ++r)
                for (int c = 0; c < 3; ++c) H[r * 3 + c] = sum[8 + r * 3 + c] / n - mp[r] * mq[c];
            kabsch_rotation(H, R);
            for (int r = 0; r < 3; ++r) tr[r] = mq[r] - (R[r * 3] * mp[0] + R[r * 3 + 1] * mp[1] + R[r * 3 + 2] * mp[2]);
        }
        double Tn[16];
        for (int r = 0; r < 3; ++r) {
            for (int c = 0; c < 4; ++c)
                Tn[r * 4 + c] = R[r * 3] * S->T[c] + R[r * 3 + 1] * S->T[4 + c] + R[r * 3 + 2] * S->T[8 + c] + (c == 3 ? tr[r] : 0.0);
        }
        Tn[12] = 0; Tn[13] = 0; Tn[14] = 0; Tn[15] = 1;
        for (int e = 0; e < 16; ++e) S->T[e] = Tn[e];
        S->updates += 1;
    } else {
        S->done = 1;
    }
    for (int e = 0; e < 16; ++e) trans_out[(size_t)b * 16 + e] = (float)S->T[e];
    if (stats_out) { stats_out[b * 2] = fitness; stats_out[b * 2 + 1] = rmse; }
    if (iters_out) iters_out[b] = S->updates;
}

struct IcpWs {
    IcpGrid* grids;
    IcpState* state;
    int* cell_end;
    double* sorted_xyz;
    int* sorted_idx;
    double* partials;
    size_t bytes;
};

IcpWs carve_icp(void* ws, int B, int N) {
    IcpWs w;
    unsigned char* p = static_cast<unsigned char*>(ws);
    size_t off = 0;
    auto take = [&](size_t bytes) { void* r = p ? p + off : nullptr; off += align_up(bytes, 256); return r; };
    const int nblk = (N + kAssocThreads - 1) / kAssocThreads;
    w.grids = static_cast<IcpGrid*>(take(sizeof(IcpGrid) * B));
    w.state = static_cast<IcpState*>(take(sizeof(IcpState) * B));
    w.cell_end = static_cast<int*>(take((size_t)B * kIcpCells * 4));
    w.sorted_xyz = static_cast<double*>(take((size_t)B * N * 24));
    w.sorted_idx = static_cast<int*>(take((size_t)B * N * 4));
    w.partials = static_cast<double*>(take((size_t)B * nblk * kSums * 8));
    w.bytes = off;
    return w;
}

}  // namespace

size_t icp_workspace(int B, int N) { return carve_icp(nullptr, B, N).bytes; }

int launch_icp_refine(const float* src, const float* tgt, float* trans, int B, int N, double max_dist, int max_iter,
                      double rel_fitness, double rel_rmse, double* stats, int32_t* iters, int32_t* corr, void* ws, size_t ws_bytes,
                      cudaStream_t st) {
    IcpWs w = carve_icp(ws, B, N);
    if (ws_bytes < w.bytes) {
        set_error("icp_refine: workspace too small (%zu < %zu)", ws_bytes, w.bytes);
        return PDSC_ERR_WORKSPACE;
    }
    const size_t smem = (size_t)kIcpCells * 4;
    cudaFuncSetAttribute(icp_grid_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    icp_grid_kernel<<<B, kGridThreads, smem, st>>>(tgt, trans, N, max_dist, w.grids, w.cell_end, w.sorted_xyz, w.sorted_idx, w.state);
    const int nblk = (N + kAssocThreads - 1) / kAssocThreads;
    for (int it = 0; it <= max_iter; ++it) {
        icp_assoc_kernel<<<dim3(nblk, B), kAssocThreads, 0, st>>>(src, N, max_dist, w.grids, w.cell_end, w.sorted_xyz, w.sorted_idx,
                                                                  w.state, w.partials, corr);
        icp_update_kernel<<<B, 32, 0, st>>>(N, nblk, max_iter, rel_fitness, rel_rmse, w.partials, w.state, trans, stats, iters);
    }
    count_launches(1 + 2 * (max_iter + 1));
    return check_launch("icp_refine");
}

}  // namespace pdsc
