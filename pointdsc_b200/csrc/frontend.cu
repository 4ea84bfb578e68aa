// nvcc-flags: -fmad=false
// Demo front end on the device: what demo_registration.py:37-44,101-104 does through open3d 0.9 and
// numpy, in float64 like both of them.  Compiled without FMA contraction so that every product and sum is rounded
// like the scalar double arithmetic of the CPU restatement (oracle/restate_frontend.py): neighbour lists are
// compared bit-exactly.
//
//   grid_*            uniform grid over the cloud: bounding box (atomics on ordered keys), cell histogram, tiled scan,
//                     scatter into cell order
//   hybrid_kernel     the max_nn nearest points with dist^2 < radius^2 (FLANN "hybrid" search), one thread per query,
//                     queries walked in cell order, cells visited in shells with an early exit; the candidate list lives
//                     in shared memory; optional fused covariance + smallest eigenvector (normals)
//   hybrid_warp_kernel  the same search with one warp per query and rank-by-counting selection (small clouds)
//   voxel_reduce      mean point / normalised summed normal per occupied voxel, ascending (ix,iy,iz), points summed in
//                     ascending input index
//   spfh / fpfh       Darboux-frame histograms (Rusu et al., ICRA 2009, PCL/open3d form)
//   nn_match_f64      argmin_j sqrt(2 - 2 a_i.b_j + 1e-6), first minimum (demo_registration.py:101-102)
#include <float.h>
#include <math.h>

#include <algorithm>

#include "common.cuh"

namespace pdsc {

namespace {

struct GridParams {
    double origin[3];
    double h;
    int dims[3];
    int stride[3];
    int ncells;
    int status;          // 0 ok, 1 grid does not fit the workspace (voxel mode), ncells = cells needed (saturated)
    int occupied;        // number of non-empty cells (written by the scan)
    int pad;
};

constexpr int kScanThreads = 1024;

__device__ __forceinline__ int cell_coord(double p, double origin, double h, int dim) {
    int c = (int)floor((p - origin) / h);
    return min(max(c, 0), dim - 1);
}

// Order-preserving map double -> uint64 (ascending doubles == ascending keys); keys of finite values are > 0.
__device__ __forceinline__ unsigned long long dkey(double v) {
    unsigned long long u = (unsigned long long)__double_as_longlong(v);
    return (u >> 63) ? ~u : (u | 0x8000000000000000ull);
}
__device__ __forceinline__ double dkey_inv(unsigned long long k) {
    unsigned long long u = (k >> 63) ? (k & 0x7fffffffffffffffull) : ~k;
    return __longlong_as_double((long long)u);
}

// bbox[0..2] = max over points of key(v), bbox[3..5] = max of ~key(v) (i.e. the minimum); zero-initialised by the memset.
__global__ void __launch_bounds__(256) grid_bbox_kernel(const double* __restrict__ pts, int n, unsigned long long* __restrict__ bbox) {
    double lo[3] = {DBL_MAX, DBL_MAX, DBL_MAX}, hi[3] = {-DBL_MAX, -DBL_MAX, -DBL_MAX};
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        for (int c = 0; c < 3; ++c) {
            double v = pts[(size_t)i * 3 + c];
            lo[c] = fmin(lo[c], v);
            hi[c] = fmax(hi[c], v);
        }
    for (int c = 0; c < 3; ++c) {
        for (int o = 16; o > 0; o >>= 1) {
            lo[c] = fmin(lo[c], __shfl_xor_sync(0xffffffffu, lo[c], o));
            hi[c] = fmax(hi[c], __shfl_xor_sync(0xffffffffu, hi[c], o));
        }
        if ((threadIdx.x & 31) == 0 && lo[c] <= hi[c]) {
            atomicMax(&bbox[c], dkey(hi[c]));
            atomicMax(&bbox[3 + c], ~dkey(lo[c]));
        }
    }
}

// mode 0: search grid (x fastest, cell grows until the grid fits); mode 1: voxel grid
// (origin = min - h/2 as open3d, z fastest so that the linear index ascends with (ix,iy,iz)).
__global__ void grid_setup_kernel(const unsigned long long* __restrict__ bbox, double h, int mode, int max_cells, GridParams* gp) {
    double lo[3], hi[3];
    for (int c = 0; c < 3; ++c) {
        hi[c] = dkey_inv(bbox[c]);
        lo[c] = dkey_inv(~bbox[3 + c]);
    }
    {
        GridParams g;
        g.status = 0;
        g.occupied = 0;
        g.pad = 0;
        g.h = h;
        for (int c = 0; c < 3; ++c) g.origin[c] = (mode == 1) ? lo[c] - h * 0.5 : lo[c];
        for (int it = 0; it < 200; ++it) {
            double cells = 1.0;
            for (int c = 0; c < 3; ++c) {
                double d = floor((hi[c] - g.origin[c]) / g.h) + 1.0;
                d = fmin(fmax(d, 1.0), 2.0e9);
                g.dims[c] = (int)d;
                cells *= d;
            }
            if (cells <= (double)max_cells) {
                g.ncells = (int)cells;
                break;
            }
            if (mode == 1) {
                g.status = 1;
                g.ncells = cells > 2.0e9 ? 2000000000 : (int)cells;
                break;
            }
            g.h *= 1.25;
        }
        if (mode == 1) {
            g.stride[0] = g.dims[1] * g.dims[2]; g.stride[1] = g.dims[2]; g.stride[2] = 1;
        } else {
            g.stride[0] = 1; g.stride[1] = g.dims[0]; g.stride[2] = g.dims[0] * g.dims[1];
        }
        *gp = g;
    }
}

__global__ void grid_count_kernel(const double* __restrict__ pts, int n, const GridParams* __restrict__ gp,
                                  int* __restrict__ cell_of, int* __restrict__ count) {
    if (gp->status) return;
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int c = 0;
    for (int a = 0; a < 3; ++a) c += cell_coord(pts[(size_t)i * 3 + a], gp->origin[a], gp->h, gp->dims[a]) * gp->stride[a];
    cell_of[i] = c;
    atomicAdd(&count[c], 1);
}

// Exclusive prefix sums over the cells, start[c] of count and rank[c] of (count > 0), in three steps: per-tile totals,
// a one-CTA scan of the totals, per-tile scans with the tile's offset.
constexpr int kTileCells = 4096;     // cells per scan tile (1024 threads x 4)

__device__ __forceinline__ int2 block_exclusive_scan(int2 v, int2* total) {
    __shared__ int2 warp_tot[32];
    __shared__ int2 all;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int2 inc = v;
    for (int o = 1; o < 32; o <<= 1) {
        int ta = __shfl_up_sync(0xffffffffu, inc.x, o), tb = __shfl_up_sync(0xffffffffu, inc.y, o);
        if (lane >= o) { inc.x += ta; inc.y += tb; }
    }
    __syncthreads();
    if (lane == 31) warp_tot[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        int2 w = warp_tot[lane];
        for (int o = 1; o < 32; o <<= 1) {
            int ta = __shfl_up_sync(0xffffffffu, w.x, o), tb = __shfl_up_sync(0xffffffffu, w.y, o);
            if (lane >= o) { w.x += ta; w.y += tb; }
        }
        warp_tot[lane] = w;
        if (lane == 31) all = w;
    }
    __syncthreads();
    int2 off = warp ? warp_tot[warp - 1] : make_int2(0, 0);
    if (total) *total = all;
    return make_int2(off.x + inc.x - v.x, off.y + inc.y - v.y);
}

__global__ void __launch_bounds__(kScanThreads) grid_tile_sum_kernel(const int* __restrict__ count, const GridParams* __restrict__ gp,
                                                                     int2* __restrict__ tile_tot) {
    if (gp->status) return;
    const int ncells = gp->ncells;
    const int base = blockIdx.x * kTileCells;
    if (base >= ncells) return;
    int2 v = make_int2(0, 0);
    for (int k = 0; k < 4; ++k) {
        const int c = base + k * kScanThreads + threadIdx.x;
        const int a = c < ncells ? count[c] : 0;
        v.x += a;
        v.y += a > 0;
    }
    int2 tot;
    block_exclusive_scan(v, &tot);
    if (threadIdx.x == 0) tile_tot[blockIdx.x] = tot;
}

__global__ void __launch_bounds__(kScanThreads) grid_tile_scan_kernel(int2* __restrict__ tile_tot, int* __restrict__ start,
                                                                      GridParams* gp, int* occupied_out) {
    __shared__ int2 carry;
    if (gp->status) {
        if (threadIdx.x == 0 && occupied_out) *occupied_out = -1;
        return;
    }
    const int ncells = gp->ncells;
    const int ntiles = (ncells + kTileCells - 1) / kTileCells;
    if (threadIdx.x == 0) carry = make_int2(0, 0);
    __syncthreads();
    for (int base = 0; base < ntiles; base += kScanThreads) {
        const int i = base + threadIdx.x;
        int2 v = i < ntiles ? tile_tot[i] : make_int2(0, 0);
        int2 tot;
        int2 ex = block_exclusive_scan(v, &tot);
        const int2 c = carry;
        if (i < ntiles) tile_tot[i] = make_int2(c.x + ex.x, c.y + ex.y);
        __syncthreads();
        if (threadIdx.x == 0) carry = make_int2(c.x + tot.x, c.y + tot.y);
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        start[ncells] = carry.x;
        gp->occupied = carry.y;
        if (occupied_out) *occupied_out = carry.y;
    }
}

__global__ void __launch_bounds__(kScanThreads) grid_tile_apply_kernel(const int* __restrict__ count, const int2* __restrict__ tile_off,
                                                                       const GridParams* __restrict__ gp, int* __restrict__ start,
                                                                       int* __restrict__ rank) {
    if (gp->status) return;
    const int ncells = gp->ncells;
    const int base = blockIdx.x * kTileCells;
    if (base >= ncells) return;
    // thread t owns the 4 consecutive cells base + 4t .. base + 4t + 3
    int a[4];
    int2 v = make_int2(0, 0);
    for (int k = 0; k < 4; ++k) {
        const int c = base + threadIdx.x * 4 + k;
        a[k] = c < ncells ? count[c] : 0;
        v.x += a[k];
        v.y += a[k] > 0;
    }
    int2 ex = block_exclusive_scan(v, nullptr);
    const int2 off = tile_off[blockIdx.x];
    int ra = off.x + ex.x, rb = off.y + ex.y;
    for (int k = 0; k < 4; ++k) {
        const int c = base + threadIdx.x * 4 + k;
        if (c < ncells) {
            start[c] = ra;
            if (rank) rank[c] = rb;
        }
        ra += a[k];
        rb += a[k] > 0;
    }
}

// Scatter into cell order.  sorted_xyz may be null (voxel mode keeps indices only).
__global__ void grid_scatter_kernel(const double* __restrict__ pts, int n, const GridParams* __restrict__ gp,
                                    const int* __restrict__ cell_of, const int* __restrict__ start, int* __restrict__ fill,
                                    int* __restrict__ sorted_idx, double* __restrict__ sorted_xyz) {
    if (gp->status) return;
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int c = cell_of[i];
    int pos = start[c] + atomicAdd(&fill[c], 1);
    sorted_idx[pos] = i;
    if (sorted_xyz) {
        sorted_xyz[(size_t)pos * 3 + 0] = pts[(size_t)i * 3 + 0];
        sorted_xyz[(size_t)pos * 3 + 1] = pts[(size_t)i * 3 + 1];
        sorted_xyz[(size_t)pos * 3 + 2] = pts[(size_t)i * 3 + 2];
    }
}

// ---- smallest eigenvector of a symmetric 3x3 (D. Eberly, "A Robust Eigensolver for 3x3 Symmetric Matrices") ----------
struct V3 { double x, y, z; };
__device__ __forceinline__ V3 cross3(V3 a, V3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
__device__ __forceinline__ double dot3(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ V3 scale3(V3 a, double s) { return {a.x * s, a.y * s, a.z * s}; }

__device__ V3 eig_vec0(const double A[6], double ev) {   // A = {a00,a01,a02,a11,a12,a22}
    V3 r0 = {A[0] - ev, A[1], A[2]}, r1 = {A[1], A[3] - ev, A[4]}, r2 = {A[2], A[4], A[5] - ev};
    V3 c01 = cross3(r0, r1), c02 = cross3(r0, r2), c12 = cross3(r1, r2);
    double d0 = dot3(c01, c01), d1 = dot3(c02, c02), d2 = dot3(c12, c12);
    double dmax = d0;
    int imax = 0;
    if (d1 > dmax) { dmax = d1; imax = 1; }
    if (d2 > dmax) { imax = 2; }
    if (imax == 0) { double s = sqrt(d0); return {c01.x / s, c01.y / s, c01.z / s}; }
    if (imax == 1) { double s = sqrt(d1); return {c02.x / s, c02.y / s, c02.z / s}; }
    double s = sqrt(d2);
    return {c12.x / s, c12.y / s, c12.z / s};
}

__device__ V3 eig_vec1(const double A[6], V3 w, double ev) {
    V3 U;
    if (fabs(w.x) > fabs(w.y)) {
        double inv = 1.0 / sqrt(w.x * w.x + w.z * w.z);
        U = {-w.z * inv, 0.0, w.x * inv};
    } else {
        double inv = 1.0 / sqrt(w.y * w.y + w.z * w.z);
        U = {0.0, w.z * inv, -w.y * inv};
    }
    V3 V = cross3(w, U);
    V3 AU = {A[0] * U.x + A[1] * U.y + A[2] * U.z, A[1] * U.x + A[3] * U.y + A[4] * U.z, A[2] * U.x + A[4] * U.y + A[5] * U.z};
    V3 AV = {A[0] * V.x + A[1] * V.y + A[2] * V.z, A[1] * V.x + A[3] * V.y + A[4] * V.z, A[2] * V.x + A[4] * V.y + A[5] * V.z};
    double m00 = dot3(U, AU) - ev, m01 = dot3(U, AV), m11 = dot3(V, AV) - ev;
    double a00 = fabs(m00), a01 = fabs(m01), a11 = fabs(m11);
    double cu, cv;   // result = cu * U - cv * V
    if (a00 >= a11) {
        if (!(fmax(a00, a01) > 0.0)) return U;
        if (a00 >= a01) { double t = m01 / m00, s = 1.0 / sqrt(1 + t * t); cu = t * s; cv = s; }
        else { double t = m00 / m01, s = 1.0 / sqrt(1 + t * t); cu = s; cv = t * s; }
    } else {
        if (!(fmax(a11, a01) > 0.0)) return U;
        if (a11 >= a01) { double t = m01 / m11, s = 1.0 / sqrt(1 + t * t); cu = s; cv = t * s; }
        else { double t = m11 / m01, s = 1.0 / sqrt(1 + t * t); cu = t * s; cv = s; }
    }
    return {cu * U.x - cv * V.x, cu * U.y - cv * V.y, cu * U.z - cv * V.z};
}

__device__ V3 smallest_eigenvector(const double C[6]) {
    double mx = fmax(fmax(fmax(C[0], C[1]), fmax(C[2], C[3])), fmax(C[4], C[5]));
    if (mx == 0.0) return {0.0, 0.0, 0.0};
    double A[6];
    for (int i = 0; i < 6; ++i) A[i] = C[i] / mx;
    double norm = A[1] * A[1] + A[2] * A[2] + A[4] * A[4];
    if (!(norm > 0.0)) {
        if (A[0] < A[3] && A[0] < A[5]) return {1.0, 0.0, 0.0};
        if (A[3] < A[0] && A[3] < A[5]) return {0.0, 1.0, 0.0};
        return {0.0, 0.0, 1.0};
    }
    double q = (A[0] + A[3] + A[5]) / 3.0;
    double b00 = A[0] - q, b11 = A[3] - q, b22 = A[5] - q;
    double p = sqrt((b00 * b00 + b11 * b11 + b22 * b22 + norm * 2.0) / 6.0);
    double c00 = b11 * b22 - A[4] * A[4];
    double c01 = A[1] * b22 - A[4] * A[2];
    double c02 = A[1] * A[4] - b11 * A[2];
    double det = (b00 * c00 - A[1] * c01 + A[2] * c02) / (p * p * p);
    double half = fmin(fmax(det * 0.5, -1.0), 1.0);
    double ang = acos(half) / 3.0;
    double beta2 = cos(ang) * 2.0;
    double beta0 = cos(ang + 2.09439510239319549) * 2.0;
    double beta1 = -(beta0 + beta2);
    double e0 = q + p * beta0, e1 = q + p * beta1, e2 = q + p * beta2;
    bool pos = half >= 0.0;
    double efirst = pos ? e2 : e0, eother = pos ? e0 : e2;
    V3 first = eig_vec0(A, efirst);
    if (efirst < e1 && efirst < eother) return first;
    V3 second = eig_vec1(A, first, e1);
    if (e1 < efirst && e1 < eother) return second;
    return pos ? cross3(second, first) : cross3(first, second);
}

// ---- hybrid search -------------------------------------------------------------------------------------------------
// Shared memory: d2s[max_nn][T+1] doubles, ids[max_nn][T+1] ints (slot-major: a thread's list is a column).
template <bool kNormals, int kSearchThreads>
__global__ void __launch_bounds__(kSearchThreads) hybrid_kernel(const double* __restrict__ pts, int n, const GridParams* __restrict__ gp,
                                                                const int* __restrict__ start, const int* __restrict__ sorted_idx,
                                                                const double* __restrict__ sorted_xyz, double radius, int max_nn,
                                                                int* __restrict__ idx_out, double* __restrict__ d2_out,
                                                                int* cnt_out, double* __restrict__ normals_out, int only_pending) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    constexpr int LD = kSearchThreads + 1;
    double* d2s = reinterpret_cast<double*>(smem_raw);
    int* ids = reinterpret_cast<int*>(d2s + (size_t)max_nn * LD);
    __shared__ int sh_cnt[kSearchThreads];
    __shared__ int sh_row[kSearchThreads];

    const int t = threadIdx.x;
    const int s = blockIdx.x * kSearchThreads + t;     // position in cell order
    int m = 0;
    int self = -1;
    bool live = s < n;
    if (live) {
        self = sorted_idx[s];
        // second pass behind hybrid_warp_kernel: only the queries it gave up on (count < 0) are still open
        if (only_pending && cnt_out[self] >= 0) { live = false; self = -1; }
    }
    if (only_pending && !__syncthreads_or(live ? 1 : 0)) return;
    if (live) {
        const double qx = sorted_xyz[(size_t)s * 3 + 0], qy = sorted_xyz[(size_t)s * 3 + 1], qz = sorted_xyz[(size_t)s * 3 + 2];
        const double r2 = radius * radius;
        const int dx = gp->dims[0], dy = gp->dims[1], dz = gp->dims[2];
        const int cx = cell_coord(qx, gp->origin[0], gp->h, dx), cy = cell_coord(qy, gp->origin[1], gp->h, dy),
                  cz = cell_coord(qz, gp->origin[2], gp->h, dz);
        // Cells are visited in shells of growing Chebyshev distance rho around the query's cell; whatever lies outside
        // shells 0..rho is farther than rho*h.
        const double h = gp->h;
        const int last = (int)ceil(radius / h * (1.0 + 1e-9));
        auto walk_shell = [&](int rho, auto&& visit) {
            for (int oz = -rho; oz <= rho; ++oz) {
                const int z = cz + oz;
                if (z < 0 || z >= dz) continue;
                for (int oy = -rho; oy <= rho; ++oy) {
                    const int y = cy + oy;
                    if (y < 0 || y >= dy) continue;
                    const int row = (z * dy + y) * dx;
                    if (max(abs(oz), abs(oy)) == rho) {
                        visit(start[row + max(cx - rho, 0)], start[row + min(cx + rho, dx - 1) + 1]);
                    } else {
                        if (cx - rho >= 0) visit(start[row + cx - rho], start[row + cx - rho + 1]);
                        if (cx + rho < dx) visit(start[row + cx + rho], start[row + cx + rho + 1]);
                    }
                }
            }
        };
        // One pass: the list is kept unsorted with its worst entry tracked; a closer candidate replaces the worst, which is
        // then searched again.  (A two-pass variant -- histogram of dist^2 over logarithmic bins first, then only the bin
        // holding the max_nn-th candidate competes -- was measured and is slower here: 1.08 vs 0.86 ms for 258k normals.)
        double wd = -1.0;    // current worst (largest) key in the list and its slot, valid once m == max_nn
        int wi = -1, wslot = 0;
        for (int rho = 0; rho <= last; ++rho) {
            walk_shell(rho, [&](int b, int e) {
                for (int c = b; c < e; ++c) {
                    const double ex = sorted_xyz[(size_t)c * 3 + 0] - qx, ey = sorted_xyz[(size_t)c * 3 + 1] - qy,
                                 ez = sorted_xyz[(size_t)c * 3 + 2] - qz;
                    const double d2 = ex * ex + ey * ey + ez * ez;
                    if (!(d2 < r2)) continue;
                    if (m < max_nn) {
                        const int id = sorted_idx[c];
                        d2s[m * LD + t] = d2;
                        ids[m * LD + t] = id;
                        if (d2 > wd || (d2 == wd && id > wi)) { wd = d2; wi = id; wslot = m; }
                        ++m;
                    } else if (d2 <= wd) {
                        const int id = sorted_idx[c];
                        if (d2 < wd || id < wi) {
                            d2s[wslot * LD + t] = d2;
                            ids[wslot * LD + t] = id;
                            wd = -1.0; wi = -1;
                            for (int k = 0; k < max_nn; ++k) {
                                const double dk = d2s[k * LD + t];
                                const int ik = ids[k * LD + t];
                                if (dk > wd || (dk == wd && ik > wi)) { wd = dk; wi = ik; wslot = k; }
                            }
                        }
                    }
                }
            });
            const double safe = rho * h * (1.0 - 1e-9);
            if (m == max_nn && wd < safe * safe) break;
        }
        // ascending (d2, index): selection sort of the column
        for (int a = 0; a < m - 1; ++a) {
            int best = a;
            double bd = d2s[a * LD + t];
            int bi = ids[a * LD + t];
            for (int k = a + 1; k < m; ++k) {
                const double dk = d2s[k * LD + t];
                const int ik = ids[k * LD + t];
                if (dk < bd || (dk == bd && ik < bi)) { bd = dk; bi = ik; best = k; }
            }
            if (best != a) {
                d2s[best * LD + t] = d2s[a * LD + t];
                ids[best * LD + t] = ids[a * LD + t];
                d2s[a * LD + t] = bd;
                ids[a * LD + t] = bi;
            }
        }
        if (cnt_out) cnt_out[self] = m;
        if (kNormals) {
            // open3d ComputeCovariance: nine running sums in list order, divided by m, E[xx] - E[x]E[x]
            double c0 = 0, c1 = 0, c2 = 0, c3 = 0, c4 = 0, c5 = 0, c6 = 0, c7 = 0, c8 = 0;
            for (int k = 0; k < m; ++k) {
                const int id = ids[k * LD + t];
                const double x = pts[(size_t)id * 3 + 0], y = pts[(size_t)id * 3 + 1], z = pts[(size_t)id * 3 + 2];
                c0 += x; c1 += y; c2 += z;
                c3 += x * x; c4 += x * y; c5 += x * z; c6 += y * y; c7 += y * z; c8 += z * z;
            }
            const double dm = (double)max(m, 1);
            c0 /= dm; c1 /= dm; c2 /= dm; c3 /= dm; c4 /= dm; c5 /= dm; c6 /= dm; c7 /= dm; c8 /= dm;
            const double C[6] = {c3 - c0 * c0, c4 - c0 * c1, c5 - c0 * c2, c6 - c1 * c1, c7 - c1 * c2, c8 - c2 * c2};
            V3 nrm = smallest_eigenvector(C);
            if (m < 3 || (nrm.x * nrm.x + nrm.y * nrm.y + nrm.z * nrm.z) == 0.0) nrm = {0.0, 0.0, 1.0};
            normals_out[(size_t)self * 3 + 0] = nrm.x;
            normals_out[(size_t)self * 3 + 1] = nrm.y;
            normals_out[(size_t)self * 3 + 2] = nrm.z;
        }
    }
    if (idx_out == nullptr) return;
    sh_cnt[t] = m;
    sh_row[t] = self;
    __syncthreads();
    // coalesced write-out: consecutive threads write consecutive slots of one query's row
    for (int e = t; e < kSearchThreads * max_nn; e += kSearchThreads) {
        const int q = e / max_nn, k = e - q * max_nn;
        const int row = sh_row[q];
        if (row < 0) continue;
        const int mq = sh_cnt[q];
        idx_out[(size_t)row * max_nn + k] = (k < mq) ? ids[k * LD + q] : -1;
        if (d2_out) d2_out[(size_t)row * max_nn + k] = (k < mq) ? d2s[k * LD + q] : (double)INFINITY;
    }
}

// ---- hybrid search, one WARP per query (small clouds) ---------------------------------------------------------------
// A down-sampled cloud has a few thousand queries: one thread per query leaves most of the machine idle.  Here the lanes of
// a warp share one query: they split every candidate range, append the candidates inside the radius to a buffer in shared
// memory, stop walking shells once max_nn buffered candidates are closer than the covered margin, and then every buffered
// candidate finds its rank by counting the candidates with a smaller (dist^2, index): ranks below max_nn are the answer,
// already in order.  A query with more than kWarpCap candidates inside the radius is left to hybrid_kernel (count = -1).
constexpr int kWarpCap = 1024;
constexpr int kWarpsPerCta = 4;

__global__ void __launch_bounds__(32 * kWarpsPerCta) hybrid_warp_kernel(int n, const GridParams* __restrict__ gp, const int* __restrict__ start,
                                                                        const int* __restrict__ sorted_idx,
                                                                        const double* __restrict__ sorted_xyz, double radius, int max_nn,
                                                                        int* __restrict__ idx_out, double* __restrict__ d2_out,
                                                                        int* __restrict__ cnt_out) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    double* cd = reinterpret_cast<double*>(smem_raw) + (size_t)w * kWarpCap;
    int* ci = reinterpret_cast<int*>(reinterpret_cast<double*>(smem_raw) + (size_t)kWarpsPerCta * kWarpCap) + (size_t)w * kWarpCap;
    __shared__ int count_s[kWarpsPerCta];
    const int s = blockIdx.x * kWarpsPerCta + w;        // position in cell order
    if (s >= n) return;
    const int self = sorted_idx[s];
    const double qx = sorted_xyz[(size_t)s * 3 + 0], qy = sorted_xyz[(size_t)s * 3 + 1], qz = sorted_xyz[(size_t)s * 3 + 2];
    const double r2 = radius * radius;
    const int dx = gp->dims[0], dy = gp->dims[1], dz = gp->dims[2];
    const int cx = cell_coord(qx, gp->origin[0], gp->h, dx), cy = cell_coord(qy, gp->origin[1], gp->h, dy),
              cz = cell_coord(qz, gp->origin[2], gp->h, dz);
    const double h = gp->h;
    const int last = (int)ceil(radius / h * (1.0 + 1e-9));
    if (lane == 0) count_s[w] = 0;
    __syncwarp();
    auto visit = [&](int b, int e) {
        for (int c = b + lane; c < e; c += 32) {
            const double ex = sorted_xyz[(size_t)c * 3 + 0] - qx, ey = sorted_xyz[(size_t)c * 3 + 1] - qy,
                         ez = sorted_xyz[(size_t)c * 3 + 2] - qz;
            const double d2 = ex * ex + ey * ey + ez * ez;
            if (d2 < r2) {
                const int pos = atomicAdd(&count_s[w], 1);
                if (pos < kWarpCap) { cd[pos] = d2; ci[pos] = sorted_idx[c]; }
            }
        }
    };
    int M = 0;
    for (int rho = 0; rho <= last; ++rho) {
        for (int oz = -rho; oz <= rho; ++oz) {
            const int z = cz + oz;
            if (z < 0 || z >= dz) continue;
            for (int oy = -rho; oy <= rho; ++oy) {
                const int y = cy + oy;
                if (y < 0 || y >= dy) continue;
                const int row = (z * dy + y) * dx;
                if (max(abs(oz), abs(oy)) == rho) {
                    visit(start[row + max(cx - rho, 0)], start[row + min(cx + rho, dx - 1) + 1]);
                } else {
                    if (cx - rho >= 0) visit(start[row + cx - rho], start[row + cx - rho + 1]);
                    if (cx + rho < dx) visit(start[row + cx + rho], start[row + cx + rho + 1]);
                }
            }
        }
        __syncwarp();
        M = count_s[w];
        if (M > kWarpCap) break;
        if (M >= max_nn) {
            const double safe = rho * h * (1.0 - 1e-9), safe2 = safe * safe;
            int inside = 0;
            for (int e = lane; e < M; e += 32) inside += cd[e] < safe2;
            inside = warp_sum(inside);
            if (inside >= max_nn) break;
        }
    }
    if (M > kWarpCap) {                 // too many candidates for the buffer: hybrid_kernel takes this query
        if (lane == 0) cnt_out[self] = -1;
        return;
    }
    const int m = min(M, max_nn);
    for (int e = lane; e < M; e += 32) {
        const double de = cd[e];
        const int ie = ci[e];
        int rank = 0;
        for (int f = 0; f < M; ++f) {
            const double df = cd[f];
            rank += (df < de) || (df == de && ci[f] < ie);
        }
        if (rank < max_nn) {
            idx_out[(size_t)self * max_nn + rank] = ie;
            if (d2_out) d2_out[(size_t)self * max_nn + rank] = de;
        }
    }
    for (int k = m + lane; k < max_nn; k += 32) {
        idx_out[(size_t)self * max_nn + k] = -1;
        if (d2_out) d2_out[(size_t)self * max_nn + k] = (double)INFINITY;
    }
    if (lane == 0) cnt_out[self] = m;
}

// ---- voxel down-sample ---------------------------------------------------------------------------------------------
__global__ void voxel_reduce_kernel(const double* __restrict__ pts, const double* __restrict__ normals,
                                    const GridParams* __restrict__ gp, const int* __restrict__ start,
                                    const int* __restrict__ rank, int* __restrict__ sorted_idx, double* __restrict__ out_pts,
                                    double* __restrict__ out_normals) {
    if (gp->status) return;
    int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= gp->ncells) return;
    const int b = start[c], e = start[c + 1];
    if (e == b) return;
    // the scatter order inside a cell is arbitrary: put the indices in ascending order (insertion sort, short lists)
    for (int i = b + 1; i < e; ++i) {
        int v = sorted_idx[i], j = i - 1;
        while (j >= b && sorted_idx[j] > v) { sorted_idx[j + 1] = sorted_idx[j]; --j; }
        sorted_idx[j + 1] = v;
    }
    double px = 0, py = 0, pz = 0, nx = 0, ny = 0, nz = 0;
    for (int i = b; i < e; ++i) {
        const int id = sorted_idx[i];
        px += pts[(size_t)id * 3 + 0]; py += pts[(size_t)id * 3 + 1]; pz += pts[(size_t)id * 3 + 2];
        if (normals) { nx += normals[(size_t)id * 3 + 0]; ny += normals[(size_t)id * 3 + 1]; nz += normals[(size_t)id * 3 + 2]; }
    }
    const double cnt = (double)(e - b);
    const int o = rank[c];
    out_pts[(size_t)o * 3 + 0] = px / cnt; out_pts[(size_t)o * 3 + 1] = py / cnt; out_pts[(size_t)o * 3 + 2] = pz / cnt;
    if (normals) {
        const double z = sqrt(nx * nx + ny * ny + nz * nz);
        if (z > 0.0) { nx /= z; ny /= z; nz /= z; }
        out_normals[(size_t)o * 3 + 0] = nx; out_normals[(size_t)o * 3 + 1] = ny; out_normals[(size_t)o * 3 + 2] = nz;
    }
}

// ---- FPFH ----------------------------------------------------------------------------------------------------------
__device__ __forceinline__ int hist_bin(double x) {
    int h = (int)floor(x);
    return min(max(h, 0), 10);
}

// One warp per point, lanes over the neighbours.  Every neighbour adds the same increment to one bin of each group, so a
// bin's value depends only on how many neighbours fell into it: the lanes count (shared-memory atomics), then each bin
// repeats the sequential additions of the increment (what a scalar loop over the neighbours produces, bit for bit).
__global__ void __launch_bounds__(128) spfh_kernel(const double* __restrict__ pts, const double* __restrict__ nrm, int n,
                                                   const int* __restrict__ idx, const int* __restrict__ cnt, int max_nn,
                                                   double* __restrict__ spfh) {
    __shared__ int bins[4][33];
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int i = blockIdx.x * 4 + w;
    bins[w][lane] = 0;
    if (lane == 0) bins[w][32] = 0;
    __syncwarp();
    const int m = i < n ? cnt[i] : 0;
    if (m > 1) {
        const V3 p1 = {pts[(size_t)i * 3], pts[(size_t)i * 3 + 1], pts[(size_t)i * 3 + 2]};
        const V3 n1 = {nrm[(size_t)i * 3], nrm[(size_t)i * 3 + 1], nrm[(size_t)i * 3 + 2]};
        for (int k = 1 + lane; k < m; k += 32) {
            const int j = idx[(size_t)i * max_nn + k];
            const V3 p2 = {pts[(size_t)j * 3], pts[(size_t)j * 3 + 1], pts[(size_t)j * 3 + 2]};
            const V3 n2 = {nrm[(size_t)j * 3], nrm[(size_t)j * 3 + 1], nrm[(size_t)j * 3 + 2]};
            V3 dp = {p2.x - p1.x, p2.y - p1.y, p2.z - p1.z};
            const double dist = sqrt(dot3(dp, dp));
            double f0 = 0.0, f1 = 0.0, f2 = 0.0;
            if (dist != 0.0) {
                const double a1 = dot3(n1, dp) / dist, a2 = dot3(n2, dp) / dist;
                V3 n1c = n1, n2c = n2;
                if (acos(fabs(a1)) > acos(fabs(a2))) {
                    n1c = n2; n2c = n1;
                    dp = {-dp.x, -dp.y, -dp.z};
                    f2 = -a2;
                } else {
                    f2 = a1;
                }
                V3 v = cross3(dp, n1c);
                const double vn = sqrt(dot3(v, v));
                if (vn != 0.0) {
                    v = {v.x / vn, v.y / vn, v.z / vn};
                    const V3 ww = cross3(n1c, v);
                    f1 = dot3(v, n2c);
                    f0 = atan2(dot3(ww, n2c), dot3(n1c, n2c));
                } else {
                    f2 = 0.0;
                }
            }
            atomicAdd(&bins[w][hist_bin(11 * (f0 + M_PI) / (2.0 * M_PI))], 1);
            atomicAdd(&bins[w][11 + hist_bin(11 * (f1 + 1.0) * 0.5)], 1);
            atomicAdd(&bins[w][22 + hist_bin(11 * (f2 + 1.0) * 0.5)], 1);
        }
    }
    __syncwarp();
    if (i < n) {
        const double incr = 100.0 / (double)max(m - 1, 1);
        for (int b = lane; b < 33; b += 32) {
            double v = 0.0;
            for (int c = bins[w][b]; c > 0; --c) v += incr;
            spfh[(size_t)i * 33 + b] = v;
        }
    }
}

// one warp per point: lane b owns bins b and b+32
__global__ void __launch_bounds__(128) fpfh_kernel(const double* __restrict__ spfh, int n, const int* __restrict__ idx,
                                                   const double* __restrict__ d2, const int* __restrict__ cnt, int max_nn,
                                                   double* __restrict__ out) {
    __shared__ double row[4][33];
    __shared__ double scale[4][3];
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int i = blockIdx.x * 4 + w;
    const bool live = i < n;
    const int m = live ? cnt[i] : 0;
    double acc0 = 0.0, acc1 = 0.0;
    // lane l keeps list entries l, l+32, l+64, l+96 (max_nn <= 128); they are handed round by shuffles
    int jj[4];
    double dd[4];
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        const int k = c * 32 + lane;
        jj[c] = (k < m) ? idx[(size_t)i * max_nn + k] : 0;
        dd[c] = (k < m) ? d2[(size_t)i * max_nn + k] : 0.0;
    }
    for (int k0 = 0; k0 < m; k0 += 4) {
        double r0[4], r1[4], dist[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int k = k0 + u;
            const int c = (k >> 5) & 3, srcl = k & 31;
            int j = 0;
            double dk = 0.0;
#pragma unroll
            for (int cc = 0; cc < 4; ++cc)
                if (cc == c) { j = __shfl_sync(0xffffffffu, jj[cc], srcl); dk = __shfl_sync(0xffffffffu, dd[cc], srcl); }
            const bool use = k >= 1 && k < m && dk != 0.0;
            dist[u] = use ? dk : 0.0;
            r0[u] = use ? spfh[(size_t)j * 33 + lane] : 0.0;
            r1[u] = (use && lane == 0) ? spfh[(size_t)j * 33 + 32] : 0.0;
        }
#pragma unroll
        for (int u = 0; u < 4; ++u)
            if (dist[u] != 0.0) {            // ascending k: the order of a scalar loop
                acc0 += r0[u] / dist[u];
                if (lane == 0) acc1 += r1[u] / dist[u];
            }
    }
    row[w][lane] = acc0;
    if (lane == 0) row[w][32] = acc1;
    __syncwarp();
    if (lane < 3) {
        double s = 0.0;
        for (int b = 0; b < 11; ++b) s = s + row[w][lane * 11 + b];
        scale[w][lane] = (s != 0.0) ? 100.0 / s : 0.0;
    }
    __syncwarp();
    if (live) {
        for (int b = lane; b < 33; b += 32) {
            double v = (m > 1) ? row[w][b] * scale[w][b / 11] + spfh[(size_t)i * 33 + b] : 0.0;
            out[(size_t)i * 33 + b] = v;
        }
    }
}

// ---- float64 descriptor matching -------------------------------------------------------------------------------------
constexpr int kMatchThreads = 128;
constexpr int kMatchTile = 32;
constexpr int kMatchSplits = 8;

// Thread = one A row held in registers; B rows broadcast from shared memory; the B range is cut into kMatchSplits
// slices (blockIdx.y) whose (min, first argmin) are combined in slice order by match_reduce_kernel.
template <int DMAX>
__global__ void __launch_bounds__(kMatchThreads) nn_match_f64_kernel(const double* __restrict__ A, const double* __restrict__ Bm, int Na,
                                                                     int Nb, int D, double* __restrict__ part_val,
                                                                     int64_t* __restrict__ part_idx) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double* tile = reinterpret_cast<double*>(smem_raw);            // [kMatchTile][D]
    const int t = threadIdx.x;
    const int i = blockIdx.x * kMatchThreads + t;
    double a[DMAX];
#pragma unroll
    for (int d = 0; d < DMAX; ++d) a[d] = (i < Na && d < D) ? A[(size_t)i * D + d] : 0.0;
    const int chunk = (Nb + kMatchSplits - 1) / kMatchSplits;
    const int jb = blockIdx.y * chunk, je = min(Nb, jb + chunk);
    double best = INFINITY;
    int64_t bi = 0;
    for (int j0 = jb; j0 < je; j0 += kMatchTile) {
        const int rows = min(kMatchTile, je - j0);
        __syncthreads();
        for (int e = t; e < rows * D; e += kMatchThreads) tile[e] = Bm[(size_t)j0 * D + e];
        __syncthreads();
        int j = 0;
        for (; j + 4 <= rows; j += 4) {
            const double *b0 = tile + j * D, *b1 = b0 + D, *b2 = b1 + D, *b3 = b2 + D;
            double acc0 = 0.0, acc1 = 0.0, acc2 = 0.0, acc3 = 0.0;
#pragma unroll
            for (int d = 0; d < DMAX; ++d)
                if (d < D) {
                    acc0 += a[d] * b0[d];
                    acc1 += a[d] * b1[d];
                    acc2 += a[d] * b2[d];
                    acc3 += a[d] * b3[d];
                }
            const double d0 = sqrt(2 - 2 * acc0 + 1e-6), d1 = sqrt(2 - 2 * acc1 + 1e-6), d2 = sqrt(2 - 2 * acc2 + 1e-6),
                         d3 = sqrt(2 - 2 * acc3 + 1e-6);
            if (d0 < best) { best = d0; bi = j0 + j; }           // NaN never wins, first minimum kept (np.argmin)
            if (d1 < best) { best = d1; bi = j0 + j + 1; }
            if (d2 < best) { best = d2; bi = j0 + j + 2; }
            if (d3 < best) { best = d3; bi = j0 + j + 3; }
        }
        for (; j < rows; ++j) {
            const double* b = tile + j * D;
            double acc = 0.0;
#pragma unroll
            for (int d = 0; d < DMAX; ++d)
                if (d < D) acc += a[d] * b[d];
            const double dist = sqrt(2 - 2 * acc + 1e-6);
            if (dist < best) { best = dist; bi = j0 + j; }
        }
    }
    if (i < Na) {
        part_val[(size_t)blockIdx.y * Na + i] = best;
        part_idx[(size_t)blockIdx.y * Na + i] = bi;
    }
}

__global__ void match_reduce_kernel(const double* __restrict__ part_val, const int64_t* __restrict__ part_idx, int Na,
                                    int64_t* __restrict__ idx_out, double* __restrict__ val_out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= Na) return;
    double best = part_val[i];
    int64_t bi = part_idx[i];
    for (int s = 1; s < kMatchSplits; ++s) {
        const double v = part_val[(size_t)s * Na + i];
        if (v < best) { best = v; bi = part_idx[(size_t)s * Na + i]; }
    }
    idx_out[i] = bi;
    if (val_out) val_out[i] = best;
}

struct GridWs {
    GridParams* gp;
    unsigned long long* bbox;
    int2* tile_tot;
    int *count, *start, *rank, *fill, *cell_of, *sorted_idx;
    double* sorted_xyz;
    size_t bytes;
};

GridWs carve(void* ws, int n, int max_cells, bool want_xyz) {
    GridWs g;
    unsigned char* p = static_cast<unsigned char*>(ws);
    size_t off = 0;
    auto take = [&](size_t bytes) { void* r = p ? p + off : nullptr; off += align_up(bytes, 256); return r; };
    g.gp = static_cast<GridParams*>(take(sizeof(GridParams)));
    g.bbox = static_cast<unsigned long long*>(take(6 * 8));          // bbox, count, fill: one memset
    g.count = static_cast<int*>(take((size_t)(max_cells + 1) * 4));
    g.fill = static_cast<int*>(take((size_t)(max_cells + 1) * 4));
    g.start = static_cast<int*>(take((size_t)(max_cells + 1) * 4));
    g.rank = static_cast<int*>(take((size_t)(max_cells + 1) * 4));
    g.tile_tot = static_cast<int2*>(take(((size_t)max_cells / kTileCells + 2) * 8));
    g.cell_of = static_cast<int*>(take((size_t)n * 4));
    g.sorted_idx = static_cast<int*>(take((size_t)n * 4));
    g.sorted_xyz = want_xyz ? static_cast<double*>(take((size_t)n * 24)) : nullptr;
    g.bytes = off;
    return g;
}

int build_grid(const double* pts, int n, double h, int mode, int max_cells, const GridWs& g, int* occupied_out, cudaStream_t st) {
    cudaMemsetAsync(g.bbox, 0, (size_t)((unsigned char*)g.start - (unsigned char*)g.bbox), st);
    const int ntiles = max_cells / kTileCells + 1;
    grid_bbox_kernel<<<std::min((n + 255) / 256, 4 * num_sms()), 256, 0, st>>>(pts, n, g.bbox);
    grid_setup_kernel<<<1, 1, 0, st>>>(g.bbox, h, mode, max_cells, g.gp);
    grid_count_kernel<<<(n + 255) / 256, 256, 0, st>>>(pts, n, g.gp, g.cell_of, g.count);
    grid_tile_sum_kernel<<<ntiles, kScanThreads, 0, st>>>(g.count, g.gp, g.tile_tot);
    grid_tile_scan_kernel<<<1, kScanThreads, 0, st>>>(g.tile_tot, g.start, g.gp, occupied_out);
    grid_tile_apply_kernel<<<ntiles, kScanThreads, 0, st>>>(g.count, g.tile_tot, g.gp, g.start, mode == 1 ? g.rank : nullptr);
    grid_scatter_kernel<<<(n + 255) / 256, 256, 0, st>>>(pts, n, g.gp, g.cell_of, g.start, g.fill, g.sorted_idx, g.sorted_xyz);
    count_launches(7);
    return check_launch("frontend grid");
}

}  // namespace

constexpr int kSearchCells = 1 << 21;
constexpr double kSearchShells = 4.0;   // search-grid cell = radius / 4 for large clouds (grown if the grid would not fit)

size_t frontend_search_workspace(int n) { return carve(nullptr, n, kSearchCells, true).bytes; }
size_t frontend_voxel_workspace(int n, int max_cells) { return carve(nullptr, n, 0, false).bytes + 17 * ((size_t)max_cells + 128); }

int launch_hybrid_search(const double* pts, int n, double radius, int max_nn, int32_t* idx, double* d2, int32_t* cnt,
                         double* normals, void* ws, size_t ws_bytes, cudaStream_t st) {
    GridWs g = carve(ws, n, kSearchCells, true);
    if (ws_bytes < g.bytes) {
        set_error("hybrid_search: workspace too small (%zu < %zu)", ws_bytes, g.bytes);
        return PDSC_ERR_WORKSPACE;
    }
    // dense scans fill the list long before the radius is reached: fine cells pay; small (down-sampled) clouds usually
    // walk all the way out: fewer, larger cells
    const double shells = n >= 65536 ? kSearchShells : 2.0;
    int rc = build_grid(pts, n, radius / shells, 0, kSearchCells, g, nullptr, st);
    if (rc) return rc;
    // small clouds (lists only): one warp per query; whatever it gives up on is finished by the thread-per-query kernel
    const bool by_warp = !normals && idx && cnt && n < 65536;
    if (by_warp) {
        const size_t wsmem = (size_t)kWarpsPerCta * kWarpCap * 12;
        cudaFuncSetAttribute(hybrid_warp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)wsmem);
        hybrid_warp_kernel<<<(n + kWarpsPerCta - 1) / kWarpsPerCta, 32 * kWarpsPerCta, wsmem, st>>>(n, g.gp, g.start, g.sorted_idx,
                                                                                              g.sorted_xyz, radius, max_nn, idx, d2, cnt);
        count_launches(1);
    }
    // small clouds: 32-thread CTAs so that every SM gets work
    const bool small = n < 128 * 2 * num_sms();
    const int threads = small ? 32 : 128;
    const size_t smem = (size_t)max_nn * 12 * (threads + 1);
    const int blocks = (n + threads - 1) / threads;
    const int pend = by_warp ? 1 : 0;
#define PDSC_HYBRID(NRM, T)                                                                                                 \
    do {                                                                                                                    \
        cudaFuncSetAttribute(hybrid_kernel<NRM, T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);                \
        hybrid_kernel<NRM, T><<<blocks, T, smem, st>>>(pts, n, g.gp, g.start, g.sorted_idx, g.sorted_xyz, radius, max_nn, idx, d2, \
                                                        cnt, normals, pend);                                               \
    } while (0)
    if (normals) { if (small) PDSC_HYBRID(true, 32); else PDSC_HYBRID(true, 128); }
    else { if (small) PDSC_HYBRID(false, 32); else PDSC_HYBRID(false, 128); }
#undef PDSC_HYBRID
    count_launches(1);
    return check_launch("hybrid_search");
}

int launch_voxel_down_sample(const double* pts, const double* normals, int n, double voxel, double* out_pts, double* out_normals,
                             int32_t* out_count, void* ws, size_t ws_bytes, cudaStream_t st) {
    // the grid takes whatever the workspace can hold
    const size_t fixed = carve(nullptr, n, 0, false).bytes;
    if (ws_bytes < fixed + 17 * 1024) {
        set_error("voxel_down_sample: workspace too small (%zu)", ws_bytes);
        return PDSC_ERR_WORKSPACE;
    }
    long long mc = (long long)((ws_bytes - fixed) / 17) - 64;
    if (mc > 2000000000LL) mc = 2000000000LL;
    const int max_cells = (int)mc;
    GridWs g = carve(ws, n, max_cells, false);
    if (g.bytes > ws_bytes) {
        set_error("voxel_down_sample: workspace too small (%zu < %zu)", ws_bytes, g.bytes);
        return PDSC_ERR_WORKSPACE;
    }
    int rc = build_grid(pts, n, voxel, 1, max_cells, g, out_count, st);
    if (rc) return rc;
    voxel_reduce_kernel<<<(max_cells + 127) / 128, 128, 0, st>>>(pts, normals, g.gp, g.start, g.rank, g.sorted_idx, out_pts, out_normals);
    count_launches(1);
    return check_launch("voxel_down_sample");
}

int launch_fpfh(const double* pts, const double* normals, int n, const int32_t* idx, const double* d2, const int32_t* cnt, int max_nn,
                double* spfh, double* out, cudaStream_t st) {
    spfh_kernel<<<(n + 3) / 4, 128, 0, st>>>(pts, normals, n, idx, cnt, max_nn, spfh);
    fpfh_kernel<<<(n + 3) / 4, 128, 0, st>>>(spfh, n, idx, d2, cnt, max_nn, out);
    count_launches(2);
    return check_launch("fpfh");
}

size_t nn_match_f64_workspace(int Na) { return (size_t)kMatchSplits * Na * 16 + 256; }

int launch_nn_match_f64(const double* A, const double* Bm, int Na, int Nb, int D, int64_t* idx, double* val, void* ws,
                        size_t ws_bytes, cudaStream_t st) {
    if (ws_bytes < nn_match_f64_workspace(Na)) {
        set_error("nn_match_f64: workspace too small (%zu < %zu)", ws_bytes, nn_match_f64_workspace(Na));
        return PDSC_ERR_WORKSPACE;
    }
    double* part_val = static_cast<double*>(ws);
    int64_t* part_idx = reinterpret_cast<int64_t*>(part_val + (size_t)kMatchSplits * Na);
    const size_t smem = (size_t)kMatchTile * D * 8;
    dim3 grid((Na + kMatchThreads - 1) / kMatchThreads, kMatchSplits);
    if (D <= 16) nn_match_f64_kernel<16><<<grid, kMatchThreads, smem, st>>>(A, Bm, Na, Nb, D, part_val, part_idx);
    else if (D <= 33) nn_match_f64_kernel<33><<<grid, kMatchThreads, smem, st>>>(A, Bm, Na, Nb, D, part_val, part_idx);
    else nn_match_f64_kernel<64><<<grid, kMatchThreads, smem, st>>>(A, Bm, Na, Nb, D, part_val, part_idx);
    match_reduce_kernel<<<(Na + 255) / 256, 256, 0, st>>>(part_val, part_idx, Na, idx, val);
    count_launches(2);
    return check_launch("nn_match_f64");
}

}  // namespace pdsc
