// pgt_dim3.cu -- the 3-D typedefs of the reference (module-irvingkirkwood.cpp:8-13, module-thermo.cpp:7-20) on the device.
//
// The reference's 3-D classes are the same templates with Dim = 3: a box periodic on all three axes (SimBox.cc:25-52),
// Aboria's search over the 27 images, 3 x 3 tensors, a 3^4 Born tensor -- and, as written there, the SAME planar
// coarse-graining normalisations (CGFuncOctic.h:16-29 ...) and the same planar z = 0 grid (irvingkirkwood/__init__.py:45-52).
// No BASELINE configuration is 3-D, so these kernels are written for parity, not for throughput: the reference's loop nest
// one thread per grid point (IrvingKirkwood.h:74-110) and one thread per particle (ThermoCalculator.h:68-90) over a 3-D
// cell list, every discrete test (<= radius, < cut-off, id order) formed exactly as the reference forms it.
#include "pgt_kernels.cuh"
#include <math_constants.h>

#define FULL_MASK 0xffffffffu

namespace {

template <bool F32>
__device__ __forceinline__ double ld3(const void* p, long long i)
{
    if (F32) return (double)__ldg(reinterpret_cast<const float*>(p) + i);
    return __ldg(reinterpret_cast<const double*>(p) + i);
}

// Aboria enforce_domain on a periodic axis (NeighbourSearchBase.h:218-224): repeated +-(high - low)
__device__ __forceinline__ bool wrap3(double& r, double low, double high, int* eflags)
{
    if (!isfinite(r)) return false;
    const double span = high - low;
    int it = 0;
    while (r < low) {
        r += span;
        if (++it > PGT_WRAP_MAX_ITERS) { if (eflags) atomicOr(eflags, PGT_EFLAG_WRAP_RUNAWAY); return false; }
    }
    while (r >= high) {
        r -= span;
        if (++it > PGT_WRAP_MAX_ITERS) { if (eflags) atomicOr(eflags, PGT_EFLAG_WRAP_RUNAWAY); return false; }
    }
    return true;
}

template <bool F32>
__device__ __forceinline__ bool load_wrapped3(const void* pos, long long i, const PgtBox& box, double& x, double& y, double& z, int* eflags)
{
    x = ld3<F32>(pos, 3 * i);
    y = ld3<F32>(pos, 3 * i + 1);
    z = ld3<F32>(pos, 3 * i + 2);
    bool alive = wrap3(x, box.lo[0], box.hi[0], eflags);
    alive = wrap3(y, box.lo[1], box.hi[1], eflags) && alive;
    alive = wrap3(z, box.lo[2], box.hi[2], eflags) && alive;
    return alive;
}

__device__ __forceinline__ int coord3(double v, double lo, double inv_cw, int nc)
{
    const int c = __double2int_rd((v - lo) * inv_cw);
    return min(nc - 1, max(0, c));
}

// ---- stage 1: histogram with arrival ranks, (scan, scatter: the 2-D launchers), per-cell sort + gather
template <bool F32>
__global__ void __launch_bounds__(256) k3_wrap_count(const void* __restrict__ pos, long long N, PgtBox box, Pgt3Cells cells,
                                                     int* __restrict__ cell_of, int* __restrict__ cell_slot, int* __restrict__ cell_count,
                                                     int* eflags)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    double x, y, z;
    int c = -1;
    if (load_wrapped3<F32>(pos, i, box, x, y, z, eflags)) {
        const int cx = coord3(x, box.lo[0], cells.inv_cw[0], cells.ncx);
        const int cy = coord3(y, box.lo[1], cells.inv_cw[1], cells.ncy);
        const int cz = coord3(z, box.lo[2], cells.inv_cw[2], cells.ncz);
        c = (cz * cells.ncy + cy) * cells.ncx + cx;
        cell_slot[i] = atomicAdd(&cell_count[c], 1);
    }
    cell_of[i] = c;
}

// one thread per cell: ids ascending (the arrival order of the atomics is not reproducible), then the sorted arrays
template <bool F32>
__global__ void __launch_bounds__(256) k3_sort_gather(const void* __restrict__ pos, const void* __restrict__ diam, const void* __restrict__ mass,
                                                      const void* __restrict__ vel, PgtBox box, int* __restrict__ sorted_orig,
                                                      const int* __restrict__ cell_start, int ncells, Pgt3FrameOut o)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ncells) return;
    const int b = cell_start[c], e = cell_start[c + 1];
    if (e - b <= 24) {
        for (int k = b + 1; k < e; ++k) {
            const int v = sorted_orig[k];
            int j = k - 1;
            while (j >= b && sorted_orig[j] > v) { sorted_orig[j + 1] = sorted_orig[j]; --j; }
            sorted_orig[j + 1] = v;
        }
    } else {
        pgt_heapsort_ids(sorted_orig + b, e - b);
    }
    for (int p = b; p < e; ++p) {
        const long long i = sorted_orig[p];
        double x, y, z;
        load_wrapped3<F32>(pos, i, box, x, y, z, nullptr);           // same operations as k3_wrap_count: same bits
        o.sx[p] = x; o.sy[p] = y; o.sz[p] = z;
        o.sd[p] = ld3<F32>(diam, i);
        if (o.sm) {
            o.sm[p] = mass ? ld3<F32>(mass, i) : 1.0;
            o.svx[p] = vel ? ld3<F32>(vel, 3 * i) : 0.0;
            o.svy[p] = vel ? ld3<F32>(vel, 3 * i + 1) : 0.0;
            o.svz[p] = vel ? ld3<F32>(vel, 3 * i + 2) : 0.0;
        }
    }
}

// ---- Aboria euclidean_search (Search.h:180-193 image loop, :438-447 candidate test; detail/Distance.h norm2:
// accum = ((0 + dx^2) + dy^2) + dz^2, inclusive <=) around an arbitrary centre, restricted to the cells that can hold a
// hit.  Unwrapped cell coordinates u in [-nc, 2 nc - 1] name (image k, cell) pairs exactly once each: k = +1 | 0 | -1.
struct Seg3 { int c0, c1, k; };

__device__ __forceinline__ int split3(int u0, int u1, int nc, Seg3 seg[3])
{
    int n = 0;
    for (int u = u0; u <= u1;) {
        int k, c0, last;
        if (u < 0) { k = +1; c0 = u + nc; last = -1; }
        else if (u >= nc) { k = -1; c0 = u - nc; last = 2 * nc - 1; }
        else { k = 0; c0 = u; last = nc - 1; }
        const int ue = min(u1, last);
        seg[n].c0 = c0; seg[n].c1 = c0 + (ue - u); seg[n].k = k;
        ++n;
        u = ue + 1;
    }
    return n;
}

__device__ __forceinline__ void range3(double v, double rad, double lo, double inv_cw, int nc, int& u0, int& u1)
{
    const double pad = 1e-9 * (rad + fabs(v) + 1.0);
    u0 = max(__double2int_rd((v - rad - pad - lo) * inv_cw), -nc);
    u1 = min(__double2int_rd((v + rad + pad - lo) * inv_cw), 2 * nc - 1);
}

// visit(q, dx, dy, dz, accum) with d = r_q - (centre + k L)
template <class F>
__device__ __forceinline__ void search3(const Pgt3Frame& fr, const PgtBox& box, const Pgt3Cells& cells, double x, double y, double z,
                                        double rad, double r2, F&& visit)
{
    int ux0, ux1, uy0, uy1, uz0, uz1;
    range3(x, rad, box.lo[0], cells.inv_cw[0], cells.ncx, ux0, ux1);
    range3(y, rad, box.lo[1], cells.inv_cw[1], cells.ncy, uy0, uy1);
    range3(z, rad, box.lo[2], cells.inv_cw[2], cells.ncz, uz0, uz1);
    Seg3 seg[3];
    const int nseg = split3(ux0, ux1, cells.ncx, seg);
    for (int uz = uz0; uz <= uz1; ++uz) {
        int kz = 0, lz = uz;
        if (uz < 0) { kz = +1; lz += cells.ncz; } else if (uz >= cells.ncz) { kz = -1; lz -= cells.ncz; }
        const double cpz = z + kz * box.L[2];
        for (int uy = uy0; uy <= uy1; ++uy) {
            int ky = 0, ly = uy;
            if (uy < 0) { ky = +1; ly += cells.ncy; } else if (uy >= cells.ncy) { ky = -1; ly -= cells.ncy; }
            const double cpy = y + ky * box.L[1];
            const int rbase = (lz * cells.ncy + ly) * cells.ncx;
            for (int s = 0; s < nseg; ++s) {
                const Seg3 sg = s == 0 ? seg[0] : (s == 1 ? seg[1] : seg[2]);
                const double cpx = x + sg.k * box.L[0];
                const int qb = __ldg(fr.cell_start + rbase + sg.c0), qe = __ldg(fr.cell_start + rbase + sg.c1 + 1);
                for (int q = qb; q < qe; ++q) {
                    const double dx = __ldg(fr.sx + q) - cpx, dy = __ldg(fr.sy + q) - cpy, dz = __ldg(fr.sz + q) - cpz;
                    double accum = dx * dx;
                    accum = accum + dy * dy;
                    accum = accum + dz * dz;
                    if (!(accum > r2)) visit(q, dx, dy, dz, accum);
                }
            }
        }
    }
}

__device__ __forceinline__ double warp_sum3(double v)
{
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) v += __shfl_xor_sync(FULL_MASK, v, d);
    return v;
}

}  // namespace

namespace {

// FixedPointCGFunc / AdaptiveCGFunc::getBondFunc (CoarseGrainFunction.h:85-88, :123-126): the integral over s in [0, 1] of
// Delta(|a + s d|^2), a = r_i - x, d = r_j - r_i, by the context's Gauss-Legendre rule (GSL's visiting order) or GSL's QAG
__device__ __forceinline__ double bond3(const Pgt3IKArgs& A, double ax, double ay, double az, double dx, double dy, double dz, int& bad)
{
    const PgtCG cg = A.cg;
    if (A.adaptive) {
        int st = 0;
        const double B = pgt_qag15([&](double t) {
            const double px = ax + t * dx, py = ay + t * dy, pz = az + t * dz;
            return pgt_cg_delta_exact(cg, px * px + py * py + pz * pz);
        }, 0.0, 1.0, A.qag_epsabs, A.qag_epsrel, A.qag_limit, &st);
        if (st) bad |= (st == PGT_QAG_CAPACITY) ? PGT_EFLAG_QAG_CAPACITY : PGT_EFLAG_QAG_FAILED;
        return B;
    }
    double B = 0.0;
    for (int k = 0; k < A.nq; ++k) {
        const double t = __ldg(A.gl_nodes + k), px = ax + t * dx, py = ay + t * dy, pz = az + t * dz;
        B += __ldg(A.gl_weights + k) * pgt_cg_delta_exact(cg, px * px + py * py + pz * pz);
    }
    return B;
}

// IrvingKirkwood::compute (IrvingKirkwood.h:74-110) with Dim = 3, one thread per grid point: the outer search (radius
// cg_rcut around x), per hit the local observables (kinetic stress, density) and the inner search (radius max_rcut around
// r_i) with VirialStress::compute_cg / ElasticVirialStress::compute_cg (ListofCommonObservables.h:72-82, :86-123) on
// r = r_i - r_j(image) (:100).  Tensors leave in the reference's flat order a + 3 b.
template <int MODEL>
__global__ void __launch_bounds__(128) k3_ik(const PgtPot pot, const Pgt3IKArgs A, int* __restrict__ eflags)
{
    const long long g = A.grid.g0 + (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= A.grid.g1) return;
    double gx, gy, gz = 0.0;
    if (A.grid.mode == 0) { gx = A.grid.xs[g / A.grid.ny]; gy = A.grid.ys[g % A.grid.ny]; }
    else { gx = A.grid.pts[3 * g]; gy = A.grid.pts[3 * g + 1]; gz = A.grid.pts[3 * g + 2]; }
    const Pgt3Frame& fr = A.fr;
    const double rc = A.max_rcut, rc2 = rc * rc;
    const bool want_v = (A.obs_mask & PGT_OBS_VIRIAL) != 0, want_k = (A.obs_mask & PGT_OBS_KINETIC) != 0 && fr.sm != nullptr;
    const bool want_e = (A.obs_mask & PGT_OBS_ELASTIC) != 0 && A.disp != nullptr;
    double vir[9], kin[9], ela[9], rho = 0.0;
#pragma unroll
    for (int k = 0; k < 9; ++k) { vir[k] = 0.0; kin[k] = 0.0; ela[k] = 0.0; }
    int bad = 0;
    search3(fr, A.box, A.cells, gx, gy, gz, A.cg.rcut, A.cg.rcutsq, [&](int p, double ax, double ay, double az, double a2) {
        const double cgval = pgt_cg_delta_exact(A.cg, a2);                       // getDeltaFunc(x, r_i) (:88)
        rho += cgval;
        if (want_k) {
            const double m = fr.sm[p], v[3] = { fr.svx[p], fr.svy[p], fr.svz[p] };
#pragma unroll
            for (int b = 0; b < 3; ++b)
#pragma unroll
                for (int a = 0; a < 3; ++a) kin[a + 3 * b] += (m * v[a] * v[b]) * cgval;
        }
        if (!want_v && !want_e) return;
        const double di = fr.sd[p];
        double ui[3] = { 0.0, 0.0, 0.0 };
        if (want_e) {
            const size_t oi = (size_t)fr.sorig[p];
            ui[0] = A.disp[3 * oi]; ui[1] = A.disp[3 * oi + 1]; ui[2] = A.disp[3 * oi + 2];
        }
        search3(fr, A.box, A.cells, fr.sx[p], fr.sy[p], fr.sz[p], rc, rc2, [&](int q, double dx, double dy, double dz, double rsq) {
            if (q == p) return;                                                  // id_i != id_j (:97)
            const double dj = __ldg(fr.sd + q);
            const double r[3] = { -dx, -dy, -dz };                               // rij = r_i - r_j (:100)
            const double f = pgt_pair_has_force<MODEL>(pot, rsq, di, dj) ? pgt_pair_eval_t<MODEL>(pot, rsq, di, dj, PGT_WANT_FDIVR) : 0.0;
            if (!want_e && f == 0.0) return;
            const double hb = 0.5 * bond3(A, ax, ay, az, dx, dy, dz, bad);        // val += 0.5 * bondval * tmp
            if (want_v && f != 0.0) {
                const double F[3] = { f * r[0], f * r[1], f * r[2] };
#pragma unroll
                for (int b = 0; b < 3; ++b)
#pragma unroll
                    for (int a = 0; a < 3; ++a) vir[a + 3 * b] += hb * (F[a] * r[b]);
            }
            if (want_e) {
                const size_t oj = (size_t)__ldg(fr.sorig + q);
                const double w[3] = { (A.disp[3 * oj] - ui[0]) + r[0], (A.disp[3 * oj + 1] - ui[1]) + r[1], (A.disp[3 * oj + 2] - ui[2]) + r[2] };
                const double wsq = w[0] * w[0] + w[1] * w[1] + w[2] * w[2];
                const double f1 = pgt_pair_eval_t<MODEL>(pot, wsq, di, dj, PGT_WANT_FDIVR);
                const double F[3] = { f * r[0], f * r[1], f * r[2] }, F1[3] = { f1 * w[0], f1 * w[1], f1 * w[2] };
#pragma unroll
                for (int b = 0; b < 3; ++b)
#pragma unroll
                    for (int a = 0; a < 3; ++a) ela[a + 3 * b] += hb * (F1[a] * w[b] - F[a] * r[b]);
            }
        });
    });
    if (bad) atomicOr(eflags, bad);
    const long long o = g - A.grid.g0;
    if (want_v && A.out_virial) for (int k = 0; k < 9; ++k) A.out_virial[9 * o + k] = vir[k];
    if ((A.obs_mask & PGT_OBS_KINETIC) && A.out_kinetic) for (int k = 0; k < 9; ++k) A.out_kinetic[9 * o + k] = kin[k];
    if ((A.obs_mask & PGT_OBS_DENSITY) && A.out_density) A.out_density[o] = rho;
    if (want_e && A.out_elastic) for (int k = 0; k < 9; ++k) A.out_elastic[9 * o + k] = ela[k];
}

// ThermoCalculator::compute (ThermoCalculator.h:68-90) with Dim = 3: local observables per particle, pair observables on
// id_i < id_j with rij = r_j(image) - r_i as the search hands it over (:84).  PART 0: virial 9 | kinetic 9 | energy 1 |
// elastic 9; PART 1..3: the Born tensor's 27 entries i + 3 (j + 3 k) of l = PART - 1 (ListofCommonObservables.h:157-184).
#define T3_THREADS 128
#define T3_NV 28
template <int MODEL>
__global__ void __launch_bounds__(T3_THREADS) k3_thermo(PgtPot pot, Pgt3Frame fr, PgtBox box, Pgt3Cells cells, double rc, unsigned obs_mask,
                                                        const double* __restrict__ disp, int part, int p_begin, int p_end,
                                                        double* __restrict__ partials)
{
    __shared__ double red[T3_THREADS / 32][T3_NV];
    double acc[T3_NV];
#pragma unroll
    for (int k = 0; k < T3_NV; ++k) acc[k] = 0.0;
    const double rc2 = rc * rc;
    const bool pairs = part > 0 || (obs_mask & (PGT_OBS_VIRIAL | PGT_OBS_ENERGY)) != 0 || ((obs_mask & PGT_OBS_ELASTIC) && disp);
    for (int p = p_begin + blockIdx.x * blockDim.x + threadIdx.x; p < p_end; p += gridDim.x * blockDim.x) {
        const double di = fr.sd[p];
        const int oi = fr.sorig[p];
        if (part == 0 && (obs_mask & PGT_OBS_KINETIC) && fr.sm) {
            const double m = fr.sm[p], v[3] = { fr.svx[p], fr.svy[p], fr.svz[p] };
#pragma unroll
            for (int b = 0; b < 3; ++b)
#pragma unroll
                for (int a = 0; a < 3; ++a) acc[9 + a + 3 * b] += m * v[a] * v[b];
        }
        if (!pairs) continue;
        double ui[3] = { 0.0, 0.0, 0.0 };
        const bool want_e = part == 0 && (obs_mask & PGT_OBS_ELASTIC) && disp;
        if (want_e) { ui[0] = disp[3 * (size_t)oi]; ui[1] = disp[3 * (size_t)oi + 1]; ui[2] = disp[3 * (size_t)oi + 2]; }
        search3(fr, box, cells, fr.sx[p], fr.sy[p], fr.sz[p], rc, rc2, [&](int q, double dx, double dy, double dz, double rsq) {
            const int oj = __ldg(fr.sorig + q);
            if (!(oi < oj)) return;
            const double dj = __ldg(fr.sd + q);
            const double r[3] = { dx, dy, dz };
            const double f = pgt_pair_eval_t<MODEL>(pot, rsq, di, dj, PGT_WANT_FDIVR);
            const double F[3] = { f * r[0], f * r[1], f * r[2] };
            if (part == 0) {
                if (obs_mask & PGT_OBS_VIRIAL) {
#pragma unroll
                    for (int b = 0; b < 3; ++b)
#pragma unroll
                        for (int a = 0; a < 3; ++a) acc[a + 3 * b] += F[a] * r[b];
                }
                if (obs_mask & PGT_OBS_ENERGY) acc[18] += pgt_pair_eval_t<MODEL>(pot, rsq, di, dj, PGT_WANT_ENERGY);
                if (want_e) {
                    const double w[3] = { (disp[3 * (size_t)oj] - ui[0]) + r[0], (disp[3 * (size_t)oj + 1] - ui[1]) + r[1],
                                          (disp[3 * (size_t)oj + 2] - ui[2]) + r[2] };
                    const double f1 = pgt_pair_eval_t<MODEL>(pot, w[0] * w[0] + w[1] * w[1] + w[2] * w[2], di, dj, PGT_WANT_FDIVR);
#pragma unroll
                    for (int b = 0; b < 3; ++b)
#pragma unroll
                        for (int a = 0; a < 3; ++a) acc[19 + a + 3 * b] += (f1 * w[a]) * w[b] - F[a] * r[b];
                }
            } else {
                const double st = pgt_pair_eval_t<MODEL>(pot, rsq, di, dj, PGT_WANT_STIFF);
                const double nrm = sqrt(rsq);
                const double nv[3] = { r[0] / nrm, r[1] / nrm, r[2] / nrm };
                const double nl = part == 1 ? nv[0] : (part == 2 ? nv[1] : nv[2]);
#pragma unroll
                for (int k = 0; k < 3; ++k)
#pragma unroll
                    for (int j = 0; j < 3; ++j)
#pragma unroll
                        for (int i = 0; i < 3; ++i) acc[i + 3 * (j + 3 * k)] += (st * r[i] + F[i]) * r[j] * nv[k] * nl;
            }
        });
    }
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < T3_NV; ++k) {
        const double s = warp_sum3(acc[k]);
        if (lane == 0) red[w][k] = s;
    }
    __syncthreads();
    if (threadIdx.x < T3_NV) {
        double s = 0.0;
        for (int ww = 0; ww < T3_THREADS / 32; ++ww) s += red[ww][threadIdx.x];
        partials[(size_t)blockIdx.x * T3_NV + threadIdx.x] = s;
    }
}

// out[dst[k]] = scale * sum over blocks of partials[b][k], k < nv; a warp per value, fixed-order tree
__global__ void __launch_bounds__(32 * T3_NV) k3_thermo_final(const double* __restrict__ partials, int nblocks, int nv, double scale,
                                                              double* __restrict__ out, int dst0)
{
    const int k = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (k >= nv) return;
    double s = 0.0;
    for (int b = lane; b < nblocks; b += 32) s += partials[(size_t)b * T3_NV + k];
    s = warp_sum3(s);
    if (lane == 0) out[dst0 + k] = s * scale;
}

}  // namespace

// ---- launch wrappers ---------------------------------------------------------------------------------------------------
void pgt3_launch_wrap_count(const void* pos, int is_f32, long long N, PgtBox box, Pgt3Cells cells, int* cell_of, int* cell_slot,
                            int* cell_count, int* eflags, cudaStream_t st, PgtLaunchCounter* lc)
{
    if (N <= 0) return;
    const int blocks = (int)((N + 255) / 256);
    if (is_f32) k3_wrap_count<true><<<blocks, 256, 0, st>>>(pos, N, box, cells, cell_of, cell_slot, cell_count, eflags);
    else k3_wrap_count<false><<<blocks, 256, 0, st>>>(pos, N, box, cells, cell_of, cell_slot, cell_count, eflags);
    lc->n++;
}

void pgt3_launch_sort_gather(const void* pos, const void* diam, const void* mass, const void* vel, int is_f32, PgtBox box, int* sorted_orig,
                             const int* cell_start, int ncells, Pgt3FrameOut out, cudaStream_t st, PgtLaunchCounter* lc)
{
    if (ncells <= 0) return;
    const int blocks = (ncells + 255) / 256;
    if (is_f32) k3_sort_gather<true><<<blocks, 256, 0, st>>>(pos, diam, mass, vel, box, sorted_orig, cell_start, ncells, out);
    else k3_sort_gather<false><<<blocks, 256, 0, st>>>(pos, diam, mass, vel, box, sorted_orig, cell_start, ncells, out);
    lc->n++;
}

cudaError_t pgt3_launch_ik(const PgtPot& pot, const Pgt3IKArgs& a, int* eflags, cudaStream_t st, PgtLaunchCounter* lc)
{
    const long long n = a.grid.g1 - a.grid.g0;
    if (n <= 0) return cudaSuccess;
    PGT_DISPATCH_MODEL(pot.model, (k3_ik<M><<<(unsigned)((n + 127) / 128), 128, 0, st>>>(pot, a, eflags)));
    lc->n++;
    return cudaGetLastError();
}

int pgt3_thermo_blocks(int n)
{
    const int b = (n + T3_THREADS - 1) / T3_THREADS;
    return b < 1 ? 1 : (b > 148 * 16 ? 148 * 16 : b);
}
int pgt3_thermo_nv() { return T3_NV; }

// out (device, 128 doubles): virial 9 | kinetic 9 | energy | born 81 | elastic 9 (at 100); all times inv_vol
cudaError_t pgt3_launch_thermo(const PgtPot& pot, Pgt3Frame fr, PgtBox box, Pgt3Cells cells, double max_rcut, unsigned obs_mask,
                               const double* disp, double inv_vol, double* partials, int blocks, double* out_dev, int p_begin, int p_end,
                               cudaStream_t st, PgtLaunchCounter* lc)
{
    cudaError_t e = cudaMemsetAsync(out_dev, 0, sizeof(double) * 128, st);
    if (e != cudaSuccess) return e;
    PGT_DISPATCH_MODEL(pot.model, (k3_thermo<M><<<blocks, T3_THREADS, 0, st>>>(pot, fr, box, cells, max_rcut, obs_mask, disp, 0, p_begin, p_end, partials)));
    k3_thermo_final<<<1, 32 * T3_NV, 0, st>>>(partials, blocks, 19, inv_vol, out_dev, 0);
    lc->n += 2;
    if ((obs_mask & PGT_OBS_ELASTIC) && disp) {
        k3_thermo_final<<<1, 32 * T3_NV, 0, st>>>(partials + 19, blocks, 9, inv_vol, out_dev, 100);
        lc->n++;
    }
    if (obs_mask & PGT_OBS_BORN) {
        for (int part = 1; part <= 3; ++part) {
            PGT_DISPATCH_MODEL(pot.model, (k3_thermo<M><<<blocks, T3_THREADS, 0, st>>>(pot, fr, box, cells, max_rcut, obs_mask, disp, part, p_begin, p_end, partials)));
            k3_thermo_final<<<1, 32 * T3_NV, 0, st>>>(partials, blocks, 27, inv_vol, out_dev, 19 + 27 * (part - 1));
            lc->n += 2;
        }
    }
    return cudaGetLastError();
}
