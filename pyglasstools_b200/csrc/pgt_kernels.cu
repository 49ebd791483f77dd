// pgt_kernels.cu -- sm_100a kernels for the IK / thermo pair-sum path (see pgt_kernels.cuh).
//
// Compiled with -fmad=false: products and sums are rounded separately unless fma() is written
// explicitly, so every discrete test below is the reference's ISO-C arithmetic.
#include "pgt_kernels.cuh"
#include <math_constants.h>
#include <limits.h>
#include <stdio.h>
#include <stdlib.h>
#include <algorithm>
#include <type_traits>

#define FULL_MASK 0xffffffffu

// The Gauss-Legendre nodes (0.5 +- 0.5 x_i on [0,1]) and weights (times (b-a)/2 = 0.5) belong to the CONTEXT: PgtIKArgs
// carries the 5-point rule by value (kernel-parameter constant bank) and a per-context device table for every other
// order, so two contexts on one device never see each other's rule.

// ------------------------------------------------------------------------------------------
// small helpers
template <bool F32>
__device__ __forceinline__ double ld_real(const void* p, long long i)
{
    if (F32) return (double)__ldg(reinterpret_cast<const float*>(p) + i);
    return __ldg(reinterpret_cast<const double*>(p) + i);
}

// Aboria enforce_domain on a periodic axis (NeighbourSearchBase.h:218-224): repeated +-(high-low).
__device__ __forceinline__ bool wrap_axis(double& r, double low, double high, int* eflags)
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

__device__ __forceinline__ int cell_coord(double v, double lo, double inv_cw, int nc)
{
    int c = __double2int_rd((v - lo) * inv_cw);
    return min(nc - 1, max(0, c));
}

// Loads, promotes and wraps particle i.  Returns false when the reference would delete it.
template <bool F32>
__device__ __forceinline__ bool load_wrapped(const void* pos, long long i, const PgtBox& box, double& x, double& y, int* eflags)
{
    const long long row = (long long)box.pstride * i;
    x = ld_real<F32>(pos, row);
    y = ld_real<F32>(pos, row + 1);
    const double z = box.pstride == 3 ? ld_real<F32>(pos, row + 2) : 0.0;
    bool alive = wrap_axis(x, box.lo[0], box.hi[0], eflags);
    alive = wrap_axis(y, box.lo[1], box.hi[1], eflags) && alive;
    // third axis is non-periodic in 2-D: out of [lo,hi) or non-finite => deleted
    if (!isfinite(z) || z < box.lo[2] || z >= box.hi[2]) alive = false;
    else if (z != 0.0 && alive && eflags) atomicOr(eflags, PGT_EFLAG_Z_NONZERO);
    return alive;
}

// ------------------------------------------------------------------------------------------
// Stage 0/1a: promote + wrap + cell key + histogram.  With a slab (grid-point sharding over GPUs: this rank owns a range
// of grid columns) a particle whose wrapped x lies outside [slab_lo, slab_hi] under every periodic image is dropped here,
// exactly like a particle outside the z slab, so that the cell list, the bond list and the gather's staging shrink with
// the shard.  The slab already includes everything the shard's tiles can reach (pgt_build_cells).
template <bool F32>
__global__ void __launch_bounds__(256) k_wrap_count(const void* __restrict__ pos, long long N, PgtBox box, PgtCells cells,
                                                    int* __restrict__ cell_of, int* __restrict__ cell_slot, int* __restrict__ cell_count,
                                                    int* eflags, int slab_on, double slab_lo, double slab_hi)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    double x, y;
    int c = -1;
    if (load_wrapped<F32>(pos, i, box, x, y, eflags)) {
        bool keep = true;
        if (slab_on) {
            const double xm = x - box.L[0], xp = x + box.L[0];
            keep = (x >= slab_lo && x <= slab_hi) || (xm >= slab_lo && xm <= slab_hi) || (xp >= slab_lo && xp <= slab_hi);
        }
        if (keep) {
            const int cx = cell_coord(x, box.lo[0], cells.inv_cwx, cells.ncx);
            const int cy = cell_coord(y, box.lo[1], cells.inv_cwy, cells.ncy);
            c = cy * cells.ncx + cx;
            // the histogram's atomic also hands out this particle's arrival rank in its cell: the scatter after the scan
            // needs no second atomic pass
            cell_slot[i] = atomicAdd(&cell_count[c], 1);
        }
    }
    cell_of[i] = c;
}

void pgt_launch_wrap_count(const void* pos, int is_f32, long long N, PgtBox box, PgtCells cells, int* cell_of, int* cell_slot,
                           int* cell_count, int* eflags, int slab_on, double slab_lo, double slab_hi, cudaStream_t st,
                           PgtLaunchCounter* lc)
{
    if (N <= 0) return;
    const int blocks = (int)((N + 255) / 256);
    if (is_f32) k_wrap_count<true><<<blocks, 256, 0, st>>>(pos, N, box, cells, cell_of, cell_slot, cell_count, eflags, slab_on, slab_lo, slab_hi);
    else k_wrap_count<false><<<blocks, 256, 0, st>>>(pos, N, box, cells, cell_of, cell_slot, cell_count, eflags, slab_on, slab_lo, slab_hi);
    lc->n++;
}

// ------------------------------------------------------------------------------------------
// Exclusive scan of int32 (three phases; in may alias out)
#define SCAN_THREADS 256
#define SCAN_ITEMS 8
#define SCAN_TILE (SCAN_THREADS * SCAN_ITEMS)

__global__ void __launch_bounds__(SCAN_THREADS) k_scan_tile(const int* in, int* out, long long n, int* tile_sums)
{
    __shared__ int warp_sums[SCAN_THREADS / 32];
    const long long base = (long long)blockIdx.x * SCAN_TILE + (long long)threadIdx.x * SCAN_ITEMS;
    int v[SCAN_ITEMS];
    int sum = 0;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) {
        v[k] = (base + k < n) ? in[base + k] : 0;
        sum += v[k];
    }
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    int inc = sum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        int t = __shfl_up_sync(FULL_MASK, inc, d);
        if (lane >= d) inc += t;
    }
    if (lane == 31) warp_sums[w] = inc;
    __syncthreads();
    if (w == 0) {
        int ws = (lane < SCAN_THREADS / 32) ? warp_sums[lane] : 0;
        int wi = ws;
#pragma unroll
        for (int d = 1; d < SCAN_THREADS / 32; d <<= 1) {
            int t = __shfl_up_sync(FULL_MASK, wi, d);
            if (lane >= d) wi += t;
        }
        if (lane < SCAN_THREADS / 32) warp_sums[lane] = wi - ws;
        if (lane == SCAN_THREADS / 32 - 1) tile_sums[blockIdx.x] = wi;
    }
    __syncthreads();
    int excl = inc - sum + warp_sums[w];
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) {
        if (base + k < n) out[base + k] = excl;
        excl += v[k];
    }
}

__global__ void __launch_bounds__(1024) k_scan_sums(int* sums, int n)
{
    __shared__ int warp_sums[32];
    __shared__ int carry_s;
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (int base = 0; base < n; base += 1024) {
        const int idx = base + threadIdx.x;
        const int v = idx < n ? sums[idx] : 0;
        int inc = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            int t = __shfl_up_sync(FULL_MASK, inc, d);
            if (lane >= d) inc += t;
        }
        if (lane == 31) warp_sums[w] = inc;
        __syncthreads();
        if (w == 0) {
            int ws = warp_sums[lane], wi = ws;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                int t = __shfl_up_sync(FULL_MASK, wi, d);
                if (lane >= d) wi += t;
            }
            warp_sums[lane] = wi - ws;
        }
        __syncthreads();
        const int carry = carry_s;
        const int excl = inc - v + warp_sums[w] + carry;
        if (idx < n) sums[idx] = excl;
        __syncthreads();
        if (threadIdx.x == 1023) carry_s = excl + v;
        __syncthreads();
    }
}

__global__ void __launch_bounds__(SCAN_THREADS) k_scan_add(int* out, long long n, const int* tile_sums)
{
    const int add = tile_sums[blockIdx.x];
    const long long base = (long long)blockIdx.x * SCAN_TILE + (long long)threadIdx.x * SCAN_ITEMS;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k)
        if (base + k < n) out[base + k] += add;
}

// Single-pass variant: one kernel, each tile's total chained to its predecessors' by a decoupled look-back.
// `state[t]`: bits 63:62 = 0 nothing yet, 1 tile total, 2 inclusive prefix; the value in the low bits, so value and flag
// travel in one 64-bit store.  Tiles are numbered by a ticket taken when the block starts: every tile a block waits for
// is already resident, so the look-back cannot wait on a block that has not been scheduled.  Sums are int32 arithmetic
// (wrap-around as in the three-phase form; the callers check the published total).
#define SCAN_TOTAL (1ull << 62)
#define SCAN_PREFIX (2ull << 62)
#define SCAN_VALUE 0xffffffffull
static_assert(SCAN_ITEMS == 8, "k_scan_chained moves 8 items per thread as two int4");

__global__ void __launch_bounds__(SCAN_THREADS) k_scan_chained(const int* in, int* out, long long n, unsigned long long* state, int* ticket)
{
    __shared__ int warp_sums[SCAN_THREADS / 32];
    __shared__ int s_tile, s_base;
    if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1);
    __syncthreads();
    const int tile = s_tile;
    const long long base = (long long)tile * SCAN_TILE + (long long)threadIdx.x * SCAN_ITEMS;
    int v[SCAN_ITEMS];
    int sum = 0;
    const bool whole = base + SCAN_ITEMS <= n;              // buffers are 16-byte aligned and base is a multiple of 8
    if (whole) {
        const int4 a = *reinterpret_cast<const int4*>(in + base), b = *reinterpret_cast<const int4*>(in + base + 4);
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
    } else {
#pragma unroll
        for (int k = 0; k < SCAN_ITEMS; ++k) v[k] = (base + k < n) ? in[base + k] : 0;
    }
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) sum += v[k];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    int inc = sum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        int t = __shfl_up_sync(FULL_MASK, inc, d);
        if (lane >= d) inc += t;
    }
    if (lane == 31) warp_sums[w] = inc;
    __syncthreads();
    if (w == 0) {
        int ws = (lane < SCAN_THREADS / 32) ? warp_sums[lane] : 0;
        int wi = ws;
#pragma unroll
        for (int d = 1; d < SCAN_THREADS / 32; d <<= 1) {
            int t = __shfl_up_sync(FULL_MASK, wi, d);
            if (lane >= d) wi += t;
        }
        if (lane < SCAN_THREADS / 32) warp_sums[lane] = wi - ws;
        const int agg = __shfl_sync(FULL_MASK, wi, SCAN_THREADS / 32 - 1);
        int excl = 0;
        if (tile == 0) {
            if (lane == 0) *reinterpret_cast<volatile unsigned long long*>(state) = SCAN_PREFIX | (unsigned)agg;
        } else {
            if (lane == 0) *reinterpret_cast<volatile unsigned long long*>(state + tile) = SCAN_TOTAL | (unsigned)agg;
            int look = tile - 1;
            while (true) {
                const int idx = look - lane;
                unsigned long long st = SCAN_PREFIX;                           // before tile 0: prefix 0
                if (idx >= 0) st = *reinterpret_cast<volatile unsigned long long*>(state + idx);
                const unsigned pref = __ballot_sync(FULL_MASK, (st >> 62) == 2);
                const int stop = pref ? __ffs(pref) - 1 : 31;                  // nearest predecessor holding a prefix
                // only the predecessors between this tile and that one have to have published; farther ones do not matter
                const unsigned need = stop >= 31 ? FULL_MASK : ((2u << stop) - 1u);
                if (__ballot_sync(FULL_MASK, (st >> 62) == 0) & need) continue;
                int x = lane <= stop ? (int)(unsigned)(st & SCAN_VALUE) : 0;
#pragma unroll
                for (int d = 16; d >= 1; d >>= 1) x += __shfl_xor_sync(FULL_MASK, x, d);
                excl += x;
                if (pref) break;
                look -= 32;
            }
            if (lane == 0) *reinterpret_cast<volatile unsigned long long*>(state + tile) = SCAN_PREFIX | (unsigned)(excl + agg);
        }
        if (lane == 0) s_base = excl;
    }
    __syncthreads();
    int excl = inc - sum + warp_sums[w] + s_base;
    int o[SCAN_ITEMS];
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) {
        o[k] = excl;
        excl += v[k];
    }
    if (whole) {
        *reinterpret_cast<int4*>(out + base) = make_int4(o[0], o[1], o[2], o[3]);
        *reinterpret_cast<int4*>(out + base + 4) = make_int4(o[4], o[5], o[6], o[7]);
    } else {
#pragma unroll
        for (int k = 0; k < SCAN_ITEMS; ++k)
            if (base + k < n) out[base + k] = o[k];
    }
}

// scratch: ticket (one 64-bit word) + one 64-bit state per tile, or the tile sums of the three-phase form
long long pgt_scan_scratch_ints(long long n) { return 2 * ((n + SCAN_TILE - 1) / SCAN_TILE + 2); }

static int scan_single_pass()
{
    static int v = -1;
    if (v < 0) {
        const char* e = getenv("PGT_SCAN_SINGLE_PASS");
        v = e ? (atoi(e) != 0) : 1;
    }
    return v;
}

void pgt_launch_scan_exclusive(const int* in, int* out, long long n, int* tile_sums, cudaStream_t st, PgtLaunchCounter* lc)
{
    if (n <= 0) return;
    const int tiles = (int)((n + SCAN_TILE - 1) / SCAN_TILE);
    if (tiles > 1 && scan_single_pass()) {
        unsigned long long* words = reinterpret_cast<unsigned long long*>(tile_sums);
        cudaMemsetAsync(words, 0, sizeof(unsigned long long) * ((size_t)tiles + 1), st);
        k_scan_chained<<<tiles, SCAN_THREADS, 0, st>>>(in, out, n, words + 1, reinterpret_cast<int*>(words));
        lc->n++;
        return;
    }
    k_scan_tile<<<tiles, SCAN_THREADS, 0, st>>>(in, out, n, tile_sums);
    lc->n++;
    if (tiles > 1) {
        k_scan_sums<<<1, 1024, 0, st>>>(tile_sums, tiles);
        k_scan_add<<<tiles, SCAN_THREADS, 0, st>>>(out, n, tile_sums);
        lc->n += 2;
    }
}

// ------------------------------------------------------------------------------------------
// Stage 1b: counting-sort scatter of particle ids, per-cell ordering, gather into sorted SoA
__global__ void __launch_bounds__(256) k_scatter_ids(const int* __restrict__ cell_of, const int* __restrict__ cell_slot, long long N,
                                                     const int* __restrict__ cell_start, int* __restrict__ sorted_orig)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    const int c = cell_of[i];
    if (c < 0) return;
    sorted_orig[__ldg(cell_start + c) + cell_slot[i]] = (int)i;      // arrival rank taken at count time (k_wrap_count)
}

void pgt_launch_scatter_ids(const int* cell_of, const int* cell_slot, long long N, const int* cell_start, int* sorted_orig,
                            cudaStream_t st, PgtLaunchCounter* lc)
{
    if (N <= 0) return;
    k_scatter_ids<<<(int)((N + 255) / 256), 256, 0, st>>>(cell_of, cell_slot, N, cell_start, sorted_orig);
    lc->n++;
}

// The scatter leaves each cell's members in arrival order; sorting them by original index makes the layout (and
// therefore every summation order downstream) reproducible run to run.  One thread per cell sorts its (usually one or
// two) members and, in the same pass, gathers them into the sorted structure of arrays: the ids are read once, the
// sorted arrays are written by consecutive threads in (nearly) consecutive places.
template <bool F32>
__global__ void __launch_bounds__(256) k_cell_sort_gather(const void* __restrict__ pos, const void* __restrict__ diam,
                                                          const void* __restrict__ mass, const void* __restrict__ vel, PgtBox box,
                                                          int* __restrict__ sorted_orig, const int* __restrict__ cell_start, int ncells,
                                                          int n_cap, double* __restrict__ sx, double* __restrict__ sy, double2* __restrict__ sxy,
                                                          double* __restrict__ sd, double* __restrict__ skxx, double* __restrict__ skxy,
                                                          double* __restrict__ skyy)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ncells) return;
    // n_cap: particles the sorted arrays hold (the live count, or last frame's plus a margin when the host did not wait
    // for this frame's: an overflow is reported through the sticky error word, never written past the arrays)
    const int b = cell_start[c], e = min(cell_start[c + 1], n_cap);
    if (e - b <= 24) {
        for (int i = b + 1; i < e; ++i) {                    // insertion sort: cells hold a handful of particles
            const int key = sorted_orig[i];
            int j = i - 1;
            while (j >= b && sorted_orig[j] > key) {
                sorted_orig[j + 1] = sorted_orig[j];
                --j;
            }
            sorted_orig[j + 1] = key;
        }
    } else {
        pgt_heapsort_ids(sorted_orig + b, e - b);            // a clustered frame or a large cell side: O(k log k), not O(k^2)
    }
    for (int p = b; p < e; ++p) {
        const long long i = sorted_orig[p];
        double x, y;
        load_wrapped<F32>(pos, i, box, x, y, nullptr);
        sx[p] = x;
        sy[p] = y;
        sxy[p] = make_double2(x, y);                         // 16-byte rows: what the quad gather's bulk copies move
        sd[p] = ld_real<F32>(diam, i);
        if (skxx) {
            const double m = mass ? ld_real<F32>(mass, i) : 1.0;
            double vx = 0.0, vy = 0.0;
            if (vel) { vx = ld_real<F32>(vel, 3 * i); vy = ld_real<F32>(vel, 3 * i + 1); }
            // KineticStress::compute (ListofCommonObservables.h:131-142): m * v[a] * v[b]
            skxx[p] = m * vx * vx;
            skxy[p] = m * vx * vy;
            skyy[p] = m * vy * vy;
        }
    }
}

void pgt_launch_cell_sort_gather(const void* pos, const void* diam, const void* mass, const void* vel, int is_f32, PgtBox box,
                                 int* sorted_orig, const int* cell_start, int ncells, int n_cap, double* sx, double* sy, double2* sxy,
                                 double* sd, double* skxx, double* skxy, double* skyy, cudaStream_t st, PgtLaunchCounter* lc)
{
    if (ncells <= 0 || n_cap <= 0) return;
    const int blocks = (ncells + 255) / 256;
    if (is_f32) k_cell_sort_gather<true><<<blocks, 256, 0, st>>>(pos, diam, mass, vel, box, sorted_orig, cell_start, ncells, n_cap, sx, sy, sxy, sd, skxx, skxy, skyy);
    else k_cell_sort_gather<false><<<blocks, 256, 0, st>>>(pos, diam, mass, vel, box, sorted_orig, cell_start, ncells, n_cap, sx, sy, sxy, sd, skxx, skxy, skyy);
    lc->n++;
}

// ------------------------------------------------------------------------------------------
// Neighbour enumeration shared by the bond and thermo kernels.
//
// Visits every sorted particle q != p whose stored position lies within `rc` (inclusive) of
// an image centre  c = r_p + k L,  k in {-1,0,1} per periodic axis -- Aboria's euclidean_search
// (Search.h:186-193, 438-447) restricted to the cells that can contain a hit.  dx = r_q - c is
// formed exactly as Aboria forms it, and accum = dx^2 + dy^2 (each product and the sum rounded
// separately; the reference's third term is +0 in 2-D).
struct Seg { int c0, c1, k; };

__device__ __forceinline__ int split_segments(int u0, int u1, int nc, Seg seg[3])
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

// visit(q, dx, dy, accum, code): code = (kx + 1) | (ky + 1) << 2 names the periodic image the hit was found under;
// bit 4 of code: q comes AFTER p in the visiting order (rows, then cells of a row in unwrapped coordinates, then sorted
// index).  The order is the same whichever end of a pair looks, mirrored, so "after" names exactly one end of every
// (pair, image): the bond's PRIMARY end is the one that sees the other after itself.  Away from the periodic seam this
// is p < q; at the seam it still splits a particle's neighbours into a leading run of secondary records and a trailing
// run of primary ones, which is what the quad gather relies on.
template <class F>
__device__ __forceinline__ void for_each_neighbor_img(const PgtFrameView& fr, const PgtBox& box, const PgtCells& cells, int p,
                                                      double xi, double yi, double rc2, F&& visit)
{
    const int cx = cell_coord(xi, box.lo[0], cells.inv_cwx, cells.ncx);
    const int cy = cell_coord(yi, box.lo[1], cells.inv_cwy, cells.ncy);
    Seg seg[3];
    const int nseg = split_segments(cx - cells.reach, cx + cells.reach, cells.ncx, seg);
    int s_home = 0;                                 // the segment that holds p's own cell (the one without an image shift)
    for (int s = 1; s < nseg; ++s) if (seg[s].k == 0) s_home = s;
    for (int uy = cy - cells.reach; uy <= cy + cells.reach; ++uy) {
        int ky = 0, row = uy;
        if (uy < 0) { ky = +1; row = uy + cells.ncy; }
        else if (uy >= cells.ncy) { ky = -1; row = uy - cells.ncy; }
        const double cpy = yi + ky * box.L[1];
        const int rbase = row * cells.ncx;
        for (int s = 0; s < nseg; ++s) {
            const double cpx = xi + seg[s].k * box.L[0];
            const int qb = __ldg(fr.cell_start + rbase + seg[s].c0);
            const int qe = __ldg(fr.cell_start + rbase + seg[s].c1 + 1);
            const int dir = uy != cy ? (uy > cy ? 1 : -1) : (s != s_home ? (s > s_home ? 1 : -1) : 0);
            const unsigned code = (unsigned)(seg[s].k + 1) | ((unsigned)(ky + 1) << 2);
            for (int q = qb; q < qe; ++q) {
                const double dx = __ldg(fr.sx + q) - cpx;
                const double dy = __ldg(fr.sy + q) - cpy;
                const double accum = dx * dx + dy * dy;
                // the particle itself is skipped (IrvingKirkwood.h:97, ThermoCalculator.h:81), tested on hits only
                if (!(accum > rc2) && q != p) visit(q, dx, dy, accum, code | ((dir > 0 || (dir == 0 && q > p)) ? 16u : 0u));
            }
        }
    }
}

template <class F>
__device__ __forceinline__ void for_each_neighbor(const PgtFrameView& fr, const PgtBox& box, const PgtCells& cells, int p,
                                                  double xi, double yi, double rc2, F&& visit)
{
    for_each_neighbor_img(fr, box, cells, p, xi, yi, rc2, [&](int q, double dx, double dy, double accum, unsigned) { visit(q, dx, dy, accum); });
}

// Stage 2/3.  MODE 0: count force-carrying half-bonds; 1: fill the CSR records;
//             MODE 2: count neighbours within max_rcut; 3: fill j only (parity export).
template <int MODEL, int MODE>
__global__ void __launch_bounds__(128) k_bonds(PgtPot pot, PgtFrameView fr, PgtBox box, PgtCells cells, double rc2,
                                               int* __restrict__ counts, PgtBondOut out, double cscale,
                                               int* __restrict__ nbr_j)
{
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= fr.n) return;
    const double xi = fr.sx[p], yi = fr.sy[p], di = fr.sd[p];
    int n = 0;
    int w = (MODE == 1 || MODE == 3) ? counts[p] : 0;
    for_each_neighbor(fr, box, cells, p, xi, yi, rc2, [&](int q, double dx, double dy, double rsq) {
        if (MODE >= 2) {
            if (MODE == 3) nbr_j[w++] = q;
            ++n;
            return;
        }
        const double dj = __ldg(fr.sd + q);
        if (!pgt_pair_has_force<MODEL>(pot, rsq, di, dj)) return;
        if (MODE == 1) {
            const double f = pgt_pair_eval_t<MODEL>(pot, rsq, di, dj, PGT_WANT_FDIVR);
            out.b[w] = make_double2(dx, dy);
            out.cf[w] = make_double2(fma(dx, dx, dy * dy) * cscale, f);
            out.j[w] = q;
            ++w;
        }
        ++n;
    });
    if (MODE == 0 || MODE == 2) counts[p] = n;
}

// Stage 2/3, single search: k_bonds_find keeps the force-carrying neighbours of every particle (index + image code,
// in the search's visiting order) in a [slot][particle] scratch list and counts them; after the scan
// k_bonds_fill_list re-forms dx exactly as the search did (same operands, same operations -> same bits) and writes
// the records.  Halves the neighbour searches of the count + fill pair above, which stays as the path for frames
// where some particle has more than PGT_NB_SLOTS bonds.
// The search of one particle: keep(slot, entry) receives the force-carrying neighbours in visiting order (entry = sorted
// index | image code << 28); returns their number, pm = the primary mask of the first PGT_NB_SLOTS of them.
template <int MODEL, int UNR, class KEEP>
__device__ __forceinline__ int bonds_search(const PgtPot& pot, const PgtFrameView& fr, const PgtBox& box, const PgtCells& cells, double rc2,
                                            int p, double xi, double yi, double di, unsigned& pm_out, KEEP&& keep)
{
    int n = 0;
    unsigned pm = 0u;                              // bit k: record k of this particle is the bond's primary end (q visited after p)
    // The neighbour enumeration of for_each_neighbor_img (same cells, same order, same operations on the same operands),
    // written as ONE flat loop over the candidates of all (row, segment) spans: the 32 searches of a warp then cost the
    // largest TOTAL candidate count among its lanes, not the sum over the spans of the largest span.
    const int cx = cell_coord(xi, box.lo[0], cells.inv_cwx, cells.ncx);
    const int cy = cell_coord(yi, box.lo[1], cells.inv_cwy, cells.ncy);
    Seg seg[3];
    const int nseg = split_segments(cx - cells.reach, cx + cells.reach, cells.ncx, seg);
    int s_home = 0;                                // the segment that holds p's own cell (the one without an image shift)
    for (int s = 1; s < nseg; ++s) if (seg[s].k == 0) s_home = s;
    int uy = cy - cells.reach, s = -1;             // current span
    const int uy_end = cy + cells.reach;
    int q = 0, qe = 0, fwd_from = 0;               // fwd_from: first q of the span that comes after p in the visiting order
    double cpx = 0.0, cpy = 0.0;
    unsigned code = 0u;
    for (;;) {
        if (q >= qe) {                             // next span (an empty one costs one turn)
            if (++s == nseg) { s = 0; ++uy; }
            if (uy > uy_end) break;
            int ky = 0, row = uy;
            if (uy < 0) { ky = +1; row = uy + cells.ncy; }
            else if (uy >= cells.ncy) { ky = -1; row = uy - cells.ncy; }
            const Seg sg = s == 0 ? seg[0] : (s == 1 ? seg[1] : seg[2]);   // selects, not a local-memory array
            cpy = yi + ky * box.L[1];
            cpx = xi + sg.k * box.L[0];
            const int rbase = row * cells.ncx;
            q = __ldg(fr.cell_start + rbase + sg.c0);
            qe = __ldg(fr.cell_start + rbase + sg.c1 + 1);
            code = (unsigned)(sg.k + 1) | ((unsigned)(ky + 1) << 2);
            const int dir = uy != cy ? (uy > cy ? 1 : -1) : (s != s_home ? (s > s_home ? 1 : -1) : 0);
            fwd_from = dir > 0 ? q : (dir < 0 ? qe : p + 1);
            continue;
        }
        // UNR candidates of the span per turn: their loads are issued together (the search is bound by the latency of
        // these loads, not by their bytes); the tests and the hits keep the visiting order
        const int m = min(qe - q, UNR);
        double rsqv[UNR], djv[UNR];
#pragma unroll
        for (int u = 0; u < UNR; ++u) {
            const int qq = u < m ? q + u : q;
            const double2 r = __ldg(fr.sxy + qq);
            djv[u] = __ldg(fr.sd + qq);                 // with the position: a hit does not wait for a second load
            const double dx = r.x - cpx, dy = r.y - cpy;
            rsqv[u] = dx * dx + dy * dy;
        }
#pragma unroll
        for (int u = 0; u < UNR; ++u) {
            const double rsq = rsqv[u];
            const int qq = q + u;
            // the particle itself is skipped (IrvingKirkwood.h:97), tested on hits only
            if (u < m && !(rsq > rc2) && qq != p) {
                const double dj = djv[u];
                if (pgt_pair_has_force<MODEL>(pot, rsq, di, dj)) {
                    if (n < PGT_NB_SLOTS) {
                        keep(n, (unsigned)qq | (code << 28));
                        if (qq >= fwd_from) pm |= 1u << n;
                    }
                    ++n;
                }
            }
        }
        q += m;
    }
    pm_out = pm;
    return n;
}

template <int MODEL, int UNR>
__global__ void __launch_bounds__(128) k_bonds_find(PgtPot pot, PgtFrameView fr, PgtBox box, PgtCells cells, double rc2,
                                                    int* __restrict__ counts, unsigned* __restrict__ nb_list,
                                                    unsigned short* __restrict__ prim_mask, int* __restrict__ eflags)
{
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= fr.n) return;
    if (fr.n_dev && p >= __ldg(fr.n_dev)) {        // fr.n is only a bound (the host did not wait for the live count)
        counts[p] = 0;
        prim_mask[p] = 0;
        return;
    }
    const double xi = fr.sx[p], yi = fr.sy[p], di = fr.sd[p];
    unsigned pm;
    const size_t stride = (size_t)fr.n;
    const int n = bonds_search<MODEL, UNR>(pot, fr, box, cells, rc2, p, xi, yi, di, pm,
                                           [&](int slot, unsigned e) { nb_list[(size_t)slot * stride + p] = e; });
    counts[p] = n;
    prim_mask[p] = (unsigned short)pm;
    if (n > PGT_NB_SLOTS) atomicOr(eflags, PGT_EFLAG_BOND_OVERFLOW);
}

template <int MODEL>
__global__ void __launch_bounds__(128) k_bonds_fill_list(PgtPot pot, PgtFrameView fr, PgtBox box, const int* __restrict__ hb_off,
                                                         const unsigned* __restrict__ nb_list, PgtBondOut out, double cscale)
{
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= fr.n) return;
    const double xi = fr.sx[p], yi = fr.sy[p], di = fr.sd[p];
    const int w0 = hb_off[p], n = hb_off[p + 1] - w0;
    for (int k = 0; k < n; ++k) {
        const unsigned e = __ldg(nb_list + (size_t)k * fr.n + p);
        const int q = (int)(e & 0x0FFFFFFFu);
        const int kx = (int)((e >> 28) & 3u) - 1, ky = (int)(e >> 30) - 1;
        const double dx = __ldg(fr.sx + q) - (xi + kx * box.L[0]);
        const double dy = __ldg(fr.sy + q) - (yi + ky * box.L[1]);
        const double rsq = dx * dx + dy * dy;
        const double f = pgt_pair_eval_t<MODEL>(pot, rsq, di, __ldg(fr.sd + q), PGT_WANT_FDIVR);
        out.b[w0 + k] = make_double2(dx, dy);
        out.cf[w0 + k] = make_double2(fma(dx, dx, dy * dy) * cscale, f);
        out.j[w0 + k] = q;
    }
}

// Same records, written by a warp that FLATTENS the half-bonds of its 32 consecutive particles: lane <-> record, so
// every lane evaluates one pair force per step whatever the particles' bond counts (thread-per-particle runs at
// ~3.6/7 lanes), two steps are in flight at once, and the 36-byte records leave as coalesced stores.  The owner of
// record e is the last lane whose first record is <= e (offsets are non-decreasing): a 5-step shuffle search.
template <int MODEL>
__global__ void __launch_bounds__(128) k_bonds_fill_flat(PgtPot pot, PgtFrameView fr, PgtBox box, const int* __restrict__ hb_off,
                                                         const unsigned* __restrict__ nb_list, PgtBondOut out, double cscale)
{
    const int lane = threadIdx.x & 31;
    const long long cap = out.cap;                 // records the arrays hold (an estimate when the host did not wait for the count)
    const int p0 = (blockIdx.x * blockDim.x + threadIdx.x) & ~31;
    if (p0 >= fr.n) return;
    const bool have = p0 + lane < fr.n;
    const int p = have ? p0 + lane : fr.n - 1;
    const double xi = fr.sx[p], yi = fr.sy[p], di = fr.sd[p];
    const int w0 = have ? hb_off[p] : hb_off[fr.n];
    const int w1 = have ? hb_off[p + 1] : w0;
    const int W0 = __shfl_sync(FULL_MASK, w0, 0), W1 = __shfl_sync(FULL_MASK, w1, 31);
    for (int base = W0; base < W1; base += 32) {
        const int e = base + lane;
        int o = 0;
#pragma unroll
        for (int step = 16; step >= 1; step >>= 1) {
            const int v = __shfl_sync(FULL_MASK, w0, o + step);       // o + step <= 31 always
            if (v <= e) o += step;
        }
        const int k = e - __shfl_sync(FULL_MASK, w0, o);
        const double xo = __shfl_sync(FULL_MASK, xi, o), yo = __shfl_sync(FULL_MASK, yi, o), dio = __shfl_sync(FULL_MASK, di, o);
        if (e < W1 && e < cap) {
            const unsigned en = __ldg(nb_list + (size_t)k * fr.n + (p0 + o));
            const int q = (int)(en & 0x0FFFFFFFu);
            const int kx = (int)((en >> 28) & 3u) - 1, ky = (int)(en >> 30) - 1;
            const double dx = __ldg(fr.sx + q) - (xo + kx * box.L[0]);
            const double dy = __ldg(fr.sy + q) - (yo + ky * box.L[1]);
            const double rsq = dx * dx + dy * dy;
            const double f = pgt_pair_eval_t<MODEL>(pot, rsq, dio, __ldg(fr.sd + q), PGT_WANT_FDIVR);
            out.b[e] = make_double2(dx, dy);
            out.cf[e] = make_double2(fma(dx, dx, dy * dy) * cscale, f);
            out.j[e] = q;
        }
    }
}

template <int MODEL>
static void launch_bonds_model(int mode, const PgtPot& pot, PgtFrameView fr, PgtBox box, PgtCells cells, double rc2, int* counts,
                               PgtBondOut out, double cscale, int* nbr_j, cudaStream_t st)
{
    const int blocks = (fr.n + 127) / 128;
    switch (mode) {
    case 4: {
        // candidates loaded per turn (developer knob; measured at cfg5 with the diameters loaded alongside: 2: 0.544, 3: 0.536, 4: 0.569 ms per bond stage; 1 without: 0.581)
        static const int unr = getenv("PGT_FIND_UNROLL") ? atoi(getenv("PGT_FIND_UNROLL")) : 3;
        unsigned* nb = reinterpret_cast<unsigned*>(nbr_j);
        unsigned short* pmk = reinterpret_cast<unsigned short*>(out.b);
        if (unr == 1) k_bonds_find<MODEL, 1><<<blocks, 128, 0, st>>>(pot, fr, box, cells, rc2, counts, nb, pmk, out.j);
        else if (unr == 2) k_bonds_find<MODEL, 2><<<blocks, 128, 0, st>>>(pot, fr, box, cells, rc2, counts, nb, pmk, out.j);
        else if (unr == 3) k_bonds_find<MODEL, 3><<<blocks, 128, 0, st>>>(pot, fr, box, cells, rc2, counts, nb, pmk, out.j);
        else k_bonds_find<MODEL, 4><<<blocks, 128, 0, st>>>(pot, fr, box, cells, rc2, counts, nb, pmk, out.j);
        break;
    }
    case 5: k_bonds_fill_flat<MODEL><<<blocks, 128, 0, st>>>(pot, fr, box, counts, reinterpret_cast<const unsigned*>(nbr_j), out, cscale); break;
    case 6: k_bonds_fill_list<MODEL><<<blocks, 128, 0, st>>>(pot, fr, box, counts, reinterpret_cast<const unsigned*>(nbr_j), out, cscale); break;
    case 0: k_bonds<MODEL, 0><<<blocks, 128, 0, st>>>(pot, fr, box, cells, rc2, counts, out, cscale, nbr_j); break;
    case 1: k_bonds<MODEL, 1><<<blocks, 128, 0, st>>>(pot, fr, box, cells, rc2, counts, out, cscale, nbr_j); break;
    case 2: k_bonds<MODEL, 2><<<blocks, 128, 0, st>>>(pot, fr, box, cells, rc2, counts, out, cscale, nbr_j); break;
    default: k_bonds<MODEL, 3><<<blocks, 128, 0, st>>>(pot, fr, box, cells, rc2, counts, out, cscale, nbr_j); break;
    }
}

void pgt_launch_bonds(int mode, const PgtPot& pot, PgtFrameView fr, PgtBox box, PgtCells cells, double max_rcut, int* counts,
                      PgtBondOut out, double cscale, int* nbr_j, cudaStream_t st, PgtLaunchCounter* lc)
{
    if (fr.n <= 0) return;
    const double rc2 = max_rcut * max_rcut;   // std::pow(max_distance, 2)
    PGT_DISPATCH_MODEL(pot.model, (launch_bonds_model<M>(mode, pot, fr, box, cells, rc2, counts, out, cscale, nbr_j, st)));
    lc->n++;
}

__global__ void __launch_bounds__(256) k_rescale_bonds(const double2* __restrict__ hb_b, double2* __restrict__ hb_cf, long long n_hb,
                                                       double cscale)
{
    const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n_hb) return;
    const double2 b = hb_b[e];
    hb_cf[e].x = fma(b.x, b.x, b.y * b.y) * cscale;
}

void pgt_launch_rescale_bonds(const double2* hb_b, double2* hb_cf, long long n_hb, double cscale, cudaStream_t st,
                              PgtLaunchCounter* lc)
{
    if (n_hb <= 0) return;
    k_rescale_bonds<<<(unsigned)((n_hb + 255) / 256), 256, 0, st>>>(hb_b, hb_cf, n_hb, cscale);
    lc->n++;
}

// ------------------------------------------------------------------------------------------
// Stage 6: thermo pair sums.  One thread per particle visits its neighbours with orig_i < orig_j
// (ThermoCalculator.h:77-86), block-reduces with warp shuffles, and a single-block second pass
// adds the per-block partials in a fixed order and multiplies by 1/vol.
#define THERMO_THREADS 128
#define THERMO_NVAL 25   // virial 4 | kinetic 4 | energy 1 | born 16

__device__ __forceinline__ double warp_sum(double v)
{
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) v += __shfl_xor_sync(FULL_MASK, v, d);
    return v;
}

template <int MODEL>
__global__ void __launch_bounds__(THERMO_THREADS) k_thermo_pairs(PgtPot pot, PgtFrameView fr, PgtBox box, PgtCells cells, double rc2,
                                                                 unsigned obs_mask, double* __restrict__ partials, int p_begin, int p_end)
{
    __shared__ double red[THERMO_THREADS / 32][THERMO_NVAL];
    double acc[THERMO_NVAL];
#pragma unroll
    for (int k = 0; k < THERMO_NVAL; ++k) acc[k] = 0.0;

    // [p_begin, p_end): this rank's share of the sorted particles (pgt_thermo_set_shard; the whole frame by default); when the
    // host did not wait for the live count (pgt_thermo_compute_async) p_end is only a bound and the count is read here
    if (fr.n_dev) p_end = min(p_end, __ldg(fr.n_dev));
    for (int p = p_begin + blockIdx.x * blockDim.x + threadIdx.x; p < p_end; p += gridDim.x * blockDim.x) {
        const double xi = fr.sx[p], yi = fr.sy[p], di = fr.sd[p];
        const int oi = fr.sorig[p];
        if ((obs_mask & PGT_OBS_KINETIC) && fr.skxx) {
            acc[4] += fr.skxx[p]; acc[5] += fr.skxy[p]; acc[6] += fr.skxy[p]; acc[7] += fr.skyy[p];
        }
        if (!(obs_mask & (PGT_OBS_VIRIAL | PGT_OBS_ENERGY | PGT_OBS_BORN))) continue;
        for_each_neighbor(fr, box, cells, p, xi, yi, rc2, [&](int q, double dx, double dy, double rsq) {
            if (!(oi < __ldg(fr.sorig + q))) return;
            const double dj = __ldg(fr.sd + q);
            const double f = pgt_pair_eval_t<MODEL>(pot, rsq, di, dj, PGT_WANT_FDIVR);
            const double Fx = f * dx, Fy = f * dy;         // rij = +dx (ThermoCalculator.h:84)
            if (obs_mask & PGT_OBS_VIRIAL) {
                acc[0] += Fx * dx; acc[1] += Fy * dx; acc[2] += Fx * dy; acc[3] += Fy * dy;   // index a + 2 b
            }
            if (obs_mask & PGT_OBS_ENERGY) acc[8] += pgt_pair_eval_t<MODEL>(pot, rsq, di, dj, PGT_WANT_ENERGY);
            if (obs_mask & PGT_OBS_BORN) {
                // (stiffness r_i + F_i) r_j n_k n_l  (ListofCommonObservables.h:157-184)
                const double st = pgt_pair_eval_t<MODEL>(pot, rsq, di, dj, PGT_WANT_STIFF);
                const double nrm = sqrt(rsq);
                const double r[2] = { dx, dy }, F[2] = { Fx, Fy }, nv[2] = { dx / nrm, dy / nrm };
#pragma unroll
                for (int l = 0; l < 2; ++l)
#pragma unroll
                    for (int k = 0; k < 2; ++k)
#pragma unroll
                        for (int j = 0; j < 2; ++j)
#pragma unroll
                            for (int i = 0; i < 2; ++i)
                                acc[9 + i + 2 * (j + 2 * (k + 2 * l))] += (st * r[i] + F[i]) * r[j] * nv[k] * nv[l];
            }
        });
    }
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < THERMO_NVAL; ++k) {
        const double s = warp_sum(acc[k]);
        if (lane == 0) red[w][k] = s;
    }
    __syncthreads();
    if (threadIdx.x < THERMO_NVAL) {
        double s = 0.0;
        for (int ww = 0; ww < THERMO_THREADS / 32; ++ww) s += red[ww][threadIdx.x];
        partials[(size_t)blockIdx.x * THERMO_NVAL + threadIdx.x] = s;
    }
}

__global__ void __launch_bounds__(32 * THERMO_NVAL) k_thermo_final(const double* __restrict__ partials, int nblocks, double inv_vol,
                                                                   double* __restrict__ out)
{
    // one warp per value; lanes stride over the blocks, then a fixed-order shuffle tree
    const int k = threadIdx.x >> 5, lane = threadIdx.x & 31;
    double s = 0.0;
    for (int b = lane; b < nblocks; b += 32) s += partials[(size_t)b * THERMO_NVAL + k];
    s = warp_sum(s);
    if (lane == 0) out[k] = s * inv_vol;      // val * (1/vol)  (ThermoProperty.h:84-87)
}

int pgt_thermo_blocks(int n)
{
    int b = (n + THERMO_THREADS - 1) / THERMO_THREADS;
    return b < 1 ? 1 : (b > 148 * 16 ? 148 * 16 : b);
}

template <int MODEL>
static void launch_thermo_model(const PgtPot& pot, PgtFrameView fr, PgtBox box, PgtCells cells, double rc2, unsigned obs_mask,
                                double* partials, int blocks, int p_begin, int p_end, cudaStream_t st)
{
    k_thermo_pairs<MODEL><<<blocks, THERMO_THREADS, 0, st>>>(pot, fr, box, cells, rc2, obs_mask, partials, p_begin, p_end);
}

void pgt_launch_thermo(const PgtPot& pot, PgtFrameView fr, PgtBox box, PgtCells cells, double max_rcut, unsigned obs_mask,
                       double inv_vol, double* partials, int max_blocks, double* out_dev, int p_begin, int p_end, cudaStream_t st,
                       PgtLaunchCounter* lc)
{
    int blocks = pgt_thermo_blocks(fr.n);
    if (blocks > max_blocks) blocks = max_blocks;
    const double rc2 = max_rcut * max_rcut;
    PGT_DISPATCH_MODEL(pot.model, (launch_thermo_model<M>(pot, fr, box, cells, rc2, obs_mask, partials, blocks, p_begin, p_end, st)));
    k_thermo_final<<<1, 32 * THERMO_NVAL, 0, st>>>(partials, blocks, inv_vol, out_dev);
    lc->n += 2;
}

// ------------------------------------------------------------------------------------------
// Stage 4/5: the IK gather.  Shared pieces first.
//
// Per (grid point x, half-bond i->j with i inside the disc of x) the squared distance along the bond, in
// units of R^2, is the quadratic  t(s) = c0 + s (c1 + s c2)  with c0 = |a|^2/R^2, c1 = 2 a.b/R^2,
// c2 = |b|^2/R^2, a = r_i - x; the coarse-graining shape is summed at the Gauss-Legendre nodes in GSL's
// order (FixedPointQuadrature.h:35-45) and  B S  (S = f b(x)b, precomputed per bond) goes into three
// accumulators; 0.5 norm sum is VirialStress::compute_cg.  No global atomics anywhere.
template <int KERNEL, int Q>
__device__ __forceinline__ double bond_quadrature(const PgtIKArgs& A, double c0, double c1, double c2)
{
    double B = 0.0;
    if (Q == 5 && KERNEL == PGT_CG_OCTIC) {
        // octic: g = 1 - t^2 is positive exactly when t < 1, so the cut-off is a clamp of g at zero, done on the sign
        // word (a cleared high word leaves at most a denormal, whose square is zero); two partial sums
        double B1 = 0.0;
#pragma unroll
        for (int k = 0; k < 5; ++k) {
            const double s = A.gl_s[k];
            const double t = fma(fma(c2, s, c1), s, c0);
            const double g = fma(-t, t, 1.0);
            const int hi = __double2hiint(g);
            const double gc = __hiloint2double(hi & ~(hi >> 31), __double2loint(g));
            if (k & 1) B1 = fma(A.gl_w[k] * gc, gc, B1); else B = fma(A.gl_w[k] * gc, gc, B);
        }
        B += B1;
    } else if (Q == 5) {
#pragma unroll
        for (int k = 0; k < 5; ++k) {
            const double s = A.gl_s[k];
            const double t = fma(fma(c2, s, c1), s, c0);
            // t < 1 on the sign/exponent word: also true for the tiny negative t rounding can produce
            if (__double2hiint(t) < 0x3FF00000) B = fma(A.gl_w[k], pgt_cg_shape<KERNEL>(t), B);
        }
    } else {
        for (int k = 0; k < A.nq; ++k) {
            const double s = __ldg(A.gl_nodes + k);
            const double t = fma(fma(c2, s, c1), s, c0);
            if (__double2hiint(t) < 0x3FF00000) B = fma(__ldg(A.gl_weights + k), pgt_cg_shape<KERNEL>(t), B);
        }
    }
    return B;
}

// Rect is discontinuous at the disc edge, so its quadrature nodes are tested in the reference's own operation order
// (CoarseGrainFunction.h:42-45, 77-88 with ri = x - dr, rij = -dx of IrvingKirkwood.h:85-101): a node can never fall on
// the other side of d^2 < R^2 than it does in the reference.  a = r_i - (x + kL) exactly as the search formed it.
template <int Q>
__device__ __forceinline__ double rect_bond_exact(const PgtIKArgs& A, double gx, double gy, double ax, double ay, double2 b)
{
    const double rix = gx - ax, riy = gy - ay;               // ri = x - dr
    const double rjx = -b.x, rjy = -b.y;                     // rij = -dx
    const double R2 = A.cg.rcutsq;
    double B = 0.0;
    const int nq = Q > 0 ? Q : A.nq;
    for (int k = 0; k < nq; ++k) {
        const double s = (Q == 5) ? A.gl_s[k] : __ldg(A.gl_nodes + k);
        const double w = (Q == 5) ? A.gl_w[k] : __ldg(A.gl_weights + k);
        const double px = rix + s * rjx, py = riy + s * rjy; // ri + s rij
        const double dx = gx - px, dy = gy - py;             // x - (ri + s rij)
        const double d2 = dx * dx + dy * dy;                 // (+ 0 for z)
        if (d2 < R2) B += w;
    }
    return B;
}

// The same sum for the quad gather's integration loop: 5-point octic with the square formed once per node.
template <int KERNEL, int Q>
__device__ __forceinline__ double ik_quadrature(const PgtIKArgs& A, double c0, double c1, double c2)
{
    if (KERNEL == PGT_CG_OCTIC && Q == 5) {
        double B0 = 0.0, B1 = 0.0;
#pragma unroll
        for (int k = 0; k < 5; ++k) {
            const double s = A.gl_s[k];
            const double t = fma(fma(c2, s, c1), s, c0);
            const double g = fma(-t, t, 1.0);
            const int hi = __double2hiint(g);
            const double gc = __hiloint2double(hi & ~(hi >> 31), __double2loint(g));
            const double gg = gc * gc;
            if (k & 1) B1 = fma(A.gl_w[k], gg, B1); else B0 = fma(A.gl_w[k], gg, B0);
        }
        return B0 + B1;
    }
    return bond_quadrature<KERNEL, Q>(A, c0, c1, c2);
}

// at most one image can be within R < L/2: the nearest one; then the exact Aboria test (Search.h:186-193, 438-447)
__device__ __forceinline__ bool in_disc_exact(double px, double py, double gx, double gy, const PgtBox& box, double R2)
{
    const double ddx = px - gx, ddy = py - gy;
    const int kx = ddx > 0.5 * box.L[0] ? +1 : (ddx < -0.5 * box.L[0] ? -1 : 0);
    const int ky = ddy > 0.5 * box.L[1] ? +1 : (ddy < -0.5 * box.L[1] ? -1 : 0);
    const double ax = px - (gx + kx * box.L[0]), ay = py - (gy + ky * box.L[1]);
    return !(ax * ax + ay * ay > R2);
}

// ------------------------------------------------------------------------------------------
// Stage 4/5, second formulation ("two-phase"): the kernel above spends ~3/4 of its issue slots on
// per-grid-point bookkeeping executed by a whole group.  Here a warp owns a patch of 32 grid points:
//
//  phase 1  thread <-> grid point.  Every thread walks the cell rows of ITS disc and tests its candidate
//           particles exactly as Aboria does (dx = r_p - (x + kL), dx^2+dy^2 <= R^2); hits are appended
//           to a per-grid-point slot list in shared memory (particle index + image code).  The loops run
//           on warp-uniform trip counts (REDUX max), so the 32 searches cost one instruction stream.
//  phase 2  half-warp <-> grid point, 2 grid points at a time.  Lanes take the slots of the grid point
//           (recomputing a = r_p - c bit-identically), prefix-sum their half-bond counts, publish
//           (2a/R^2, |a|^2/R^2) and an owner table, and then stride over the flattened half-bond list:
//           t(s) = c0 + s (c1 + s c2), Gauss-Legendre sum of the shape, B f b(x)b into 3 registers.
//           A 16-lane shuffle tree reduces; the owning thread adds the result to its totals.
//  store    thread <-> grid point again: one 32-byte store per point.
// All loop bounds in phase 2 are warp-uniform (max over the two halves) so the halves stay converged.
#define IK2_WARPS 4
#define IK2_THREADS (IK2_WARPS * 32)
#define IK2_TILE_I (IK2_WARPS * 2)
#define IK2_TILE_J 16
#define IK2_MAXS 24            // slots buffered per grid point per pass (shared memory per CTA sets the L1 size: see DESIGN 4.1)
#define IK2_OWNER_CAP 128      // owner-table entries per half-warp window

struct IK2WarpSmem {
    unsigned list[IK2_MAXS * 33];          // [slot][grid point], row padded to 33 words
    double2 s_a[2][16];                    // 2 a / R^2
    double s_c0[2][16];                    // |a|^2 / R^2
    int s_rel[2][16];                      // first half-bond of the slot minus its queue offset
    unsigned char owner[2][IK2_OWNER_CAP];
};

template <int KERNEL, int Q, bool LOCAL>
__global__ void __launch_bounds__(IK2_THREADS) k_ik_gather2(const PgtIKArgs A, const int tile_i_begin)
{
    __shared__ IK2WarpSmem smem_all[IK2_WARPS];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, half = lane >> 4, l16 = lane & 15;
    const unsigned hmask = half ? 0xffff0000u : 0x0000ffffu;
    IK2WarpSmem& sm = smem_all[warp];

    const PgtFrameView& fr = A.fr;
    const double R = A.cg.rcut, R2 = A.cg.rcutsq, iR2 = A.inv_rcutsq;
    const double Lx = A.box.L[0], Ly = A.box.L[1];
    const int ncx = A.cells.ncx, ncy = A.cells.ncy;

    // ---- my grid point
    bool valid;
    long long g = 0;
    double gx = 0.0, gy = 0.0;
    if (A.grid.mode == 0) {
        const int tiles_j = (A.grid.ny + IK2_TILE_J - 1) / IK2_TILE_J;
        const int ti = tile_i_begin + blockIdx.x / tiles_j, tj = blockIdx.x % tiles_j;
        const int i = ti * IK2_TILE_I + warp * 2 + half, j = tj * IK2_TILE_J + l16;
        valid = i < A.grid.nx && j < A.grid.ny;
        if (valid) {
            g = (long long)i * A.grid.ny + j;
            valid = g >= A.grid.g0 && g < A.grid.g1;
            gx = __ldg(A.grid.xs + i);
            gy = __ldg(A.grid.ys + j);
        }
    } else {
        g = A.grid.g0 + (long long)blockIdx.x * IK2_THREADS + tid;
        valid = g < A.grid.g1;
        if (valid) { gx = __ldg(A.grid.pts + 3 * g); gy = __ldg(A.grid.pts + 3 * g + 1); }
    }

    // ---- my cell window (unwrapped cell coordinates; k = +1 left of the box, -1 right of it)
    int uy0 = 0, nrows = 0, nseg = 0;
    Seg seg[3];
    if (valid) {
        const double padx = 1e-9 * (R + fabs(gx) + 1.0), pady = 1e-9 * (R + fabs(gy) + 1.0);
        int ux0 = __double2int_rd((gx - R - padx - A.box.lo[0]) * A.cells.inv_cwx);
        int ux1 = __double2int_rd((gx + R + padx - A.box.lo[0]) * A.cells.inv_cwx);
        int uy1;
        uy0 = __double2int_rd((gy - R - pady - A.box.lo[1]) * A.cells.inv_cwy);
        uy1 = __double2int_rd((gy + R + pady - A.box.lo[1]) * A.cells.inv_cwy);
        ux0 = max(ux0, -ncx); ux1 = min(ux1, 2 * ncx - 1);
        uy0 = max(uy0, -ncy); uy1 = min(uy1, 2 * ncy - 1);
        nseg = split_segments(ux0, ux1, ncx, seg);
        nrows = nseg > 0 ? max(0, uy1 - uy0 + 1) : 0;
    }
    const int nrows_w = __reduce_max_sync(FULL_MASK, nrows);
    const int nseg_w = __reduce_max_sync(FULL_MASK, nseg);

    double txx = 0.0, txy = 0.0, tyy = 0.0;                 // totals of MY grid point
    double tkxx = 0.0, tkxy = 0.0, tkyy = 0.0, trho = 0.0;

    int r = 0, s = 0;                                       // warp-uniform cursor over (row, segment)
    bool range_loaded = false, finished = (nrows_w == 0 || nseg_w == 0);
    int q = 0, qe = 0;                                      // my range within the current (row, segment)
    unsigned kcode = 0;
    double cpx = 0.0, cpy = 0.0;

    while (!finished) {
        // ================= phase 1: thread <-> grid point, collect in-disc particles =================
        int ns = 0;
        bool flush = false;
        while (!flush) {
            if (!range_loaded) {
                if (r >= nrows_w) { finished = true; break; }
                q = qe = 0;
                if (r < nrows && s < nseg) {
                    int uy = uy0 + r, ky = 0;
                    if (uy < 0) { ky = +1; uy += ncy; }
                    else if (uy >= ncy) { ky = -1; uy -= ncy; }
                    Seg sg = seg[0];                         // select, not index: keeps seg[] in registers
                    if (s == 1) sg = seg[1];
                    if (s == 2) sg = seg[2];
                    const int rbase = uy * ncx;
                    q = __ldg(fr.cell_start + rbase + sg.c0);
                    qe = __ldg(fr.cell_start + rbase + sg.c1 + 1);
                    cpx = gx + sg.k * Lx;                    // x + k (bmax - bmin), Search.h:186-189
                    cpy = gy + ky * Ly;
                    kcode = ((unsigned)(sg.k + 1) << 28) | ((unsigned)(ky + 1) << 30);
                }
                range_loaded = true;
            }
            const int take = min(qe - q, IK2_MAXS - ns);
            const int trip = __reduce_max_sync(FULL_MASK, take);
            for (int it = 0; it < trip; ++it) {
                if (it < take) {
                    const int p = q + it;
                    const double ax = __ldg(fr.sx + p) - cpx;
                    const double ay = __ldg(fr.sy + p) - cpy;
                    const double d2 = ax * ax + ay * ay;
                    if (!(d2 > R2)) {
                        sm.list[ns * 33 + lane] = (unsigned)p | kcode;
                        ++ns;
                    }
                }
            }
            q += take;
            if (__any_sync(FULL_MASK, q < qe)) flush = true;          // some list is full: drain first
            else {
                range_loaded = false;
                if (++s >= nseg_w) { s = 0; ++r; }
            }
        }
        __syncwarp();

        // ================= phase 2: half-warp <-> grid point =================
        if (__any_sync(FULL_MASK, ns > 0)) {
            for (int pp = 0; pp < 16; ++pp) {
                const int src = half * 16 + pp;                       // lane owning this half's grid point
                const int ns_g = __shfl_sync(FULL_MASK, ns, src);
                const int ns_max = __reduce_max_sync(FULL_MASK, ns_g);
                if (ns_max == 0) continue;
                const double gxg = __shfl_sync(FULL_MASK, gx, src), gyg = __shfl_sync(FULL_MASK, gy, src);
                double axx = 0.0, axy = 0.0, ayy = 0.0, kxx = 0.0, kxy = 0.0, kyy = 0.0, rho = 0.0;

                for (int sr = 0; sr * 16 < ns_max; ++sr) {
                    const int slot = sr * 16 + l16;
                    const bool have = slot < ns_g;
                    int nb = 0, hb0 = 0;
                    double a2x = 0.0, a2y = 0.0, c0 = 0.0;
                    if (have) {
                        const unsigned e = sm.list[slot * 33 + src];
                        const int p = (int)(e & 0x0FFFFFFFu);
                        const int kx = (int)((e >> 28) & 3u) - 1, ky = (int)(e >> 30) - 1;
                        const double ax = __ldg(fr.sx + p) - (gxg + kx * Lx);
                        const double ay = __ldg(fr.sy + p) - (gyg + ky * Ly);
                        const double d2 = ax * ax + ay * ay;
                        if (LOCAL) {
                            const double cgval = pgt_cg_delta_exact(A.cg, d2);
                            rho += cgval;
                            if (fr.skxx) {
                                kxx = fma(__ldg(fr.skxx + p), cgval, kxx);
                                kxy = fma(__ldg(fr.skxy + p), cgval, kxy);
                                kyy = fma(__ldg(fr.skyy + p), cgval, kyy);
                            }
                        }
                        if (A.obs_mask & PGT_OBS_VIRIAL) {
                            hb0 = __ldg(fr.hb_off + p);
                            nb = __ldg(fr.hb_off + p + 1) - hb0;
                        }
                        if (KERNEL == PGT_CG_RECT) { a2x = ax; a2y = ay; }      // rect: the exact-order node test wants a itself
                        else { a2x = ax * (2.0 * iR2); a2y = ay * (2.0 * iR2); }
                        c0 = d2 * iR2;
                    }
                    // prefix sum of the half-bond counts inside each half
                    int tin = nb;
#pragma unroll
                    for (int d = 1; d < 16; d <<= 1) {
                        const int t = __shfl_up_sync(FULL_MASK, tin, d, 16);
                        if (l16 >= d) tin += t;
                    }
                    const int T_g = __shfl_sync(FULL_MASK, tin, 15, 16);
                    const int off = tin - nb;
                    const int T_max = __reduce_max_sync(FULL_MASK, T_g);
                    if (T_max == 0) continue;
                    const int nb_max = __reduce_max_sync(FULL_MASK, nb);
                    __syncwarp();                                      // previous consumers are done with the tables
                    sm.s_a[half][l16] = make_double2(a2x, a2y);
                    sm.s_c0[half][l16] = c0;
                    sm.s_rel[half][l16] = hb0 - off;

                    for (int T0 = 0; T0 < T_max; T0 += IK2_OWNER_CAP) {
                        if (T0 > 0) __syncwarp();
                        for (int k = 0; k < nb_max; ++k) {
                            const int idx = off + k - T0;
                            if (k < nb && idx >= 0 && idx < IK2_OWNER_CAP) sm.owner[half][idx] = (unsigned char)l16;
                        }
                        __syncwarp();
                        const int Tw_g = min(IK2_OWNER_CAP, T_g - T0), Tw_max = min(IK2_OWNER_CAP, T_max - T0);
                        for (int t0 = 0; t0 < Tw_max; t0 += 16) {
                            const int t = t0 + l16;
                            if (t < Tw_g) {
                                const int o = sm.owner[half][t];
                                const int rec = sm.s_rel[half][o] + T0 + t;
                                const double2 b = __ldg(fr.hb_b + rec);
                                const double2 cf = __ldg(fr.hb_cf + rec);
                                const double2 a2 = sm.s_a[half][o];
                                double B;
                                if (KERNEL == PGT_CG_RECT) B = rect_bond_exact<Q>(A, gxg, gyg, a2.x, a2.y, b);
                                else {
                                    const double c1 = fma(a2.x, b.x, a2.y * b.y);
                                    B = bond_quadrature<KERNEL, Q>(A, sm.s_c0[half][o], c1, cf.x);
                                }
                                const double wf = B * cf.y;
                                const double tx = wf * b.x, ty = wf * b.y;
                                axx = fma(tx, b.x, axx);
                                axy = fma(tx, b.y, axy);
                                ayy = fma(ty, b.y, ayy);
                            }
                        }
                    }
                }
                // 16-lane reduction; the result lands in every lane of the half, lane `src` keeps it
#pragma unroll
                for (int d = 8; d >= 1; d >>= 1) {
                    axx += __shfl_xor_sync(FULL_MASK, axx, d, 16);
                    axy += __shfl_xor_sync(FULL_MASK, axy, d, 16);
                    ayy += __shfl_xor_sync(FULL_MASK, ayy, d, 16);
                    if (LOCAL) {
                        kxx += __shfl_xor_sync(FULL_MASK, kxx, d, 16);
                        kxy += __shfl_xor_sync(FULL_MASK, kxy, d, 16);
                        kyy += __shfl_xor_sync(FULL_MASK, kyy, d, 16);
                        rho += __shfl_xor_sync(FULL_MASK, rho, d, 16);
                    }
                }
                if (lane == src) {
                    txx += axx; txy += axy; tyy += ayy;
                    if (LOCAL) { tkxx += kxx; tkxy += kxy; tkyy += kyy; trho += rho; }
                }
            }
        }
        __syncwarp();
    }

    // ---- store: thread <-> grid point
    if (valid) {
        const long long lg = g - A.grid.g0;
        if (A.out_virial && (A.obs_mask & PGT_OBS_VIRIAL)) {
            const double scale = 0.5 * A.cg.norm;               // val += 0.5 * bondval * (f rij (x) rij)
            *reinterpret_cast<double4*>(A.out_virial + 4 * lg) = make_double4(scale * txx, scale * txy, scale * txy, scale * tyy);
        }
        if (LOCAL && A.out_kinetic && (A.obs_mask & PGT_OBS_KINETIC))
            *reinterpret_cast<double4*>(A.out_kinetic + 4 * lg) = make_double4(tkxx, tkxy, tkxy, tkyy);
        if (LOCAL && A.out_density && (A.obs_mask & PGT_OBS_DENSITY)) A.out_density[lg] = trho;
    }
    (void)hmask;
}

template <int KERNEL, int Q>
static cudaError_t launch_ik2_t(const PgtIKArgs& a, cudaStream_t st)
{
    long long blocks;
    int tile_i_begin = 0;
    if (a.grid.mode == 0) {
        if (a.grid.g1 <= a.grid.g0) return cudaSuccess;
        const int tiles_j = (a.grid.ny + IK2_TILE_J - 1) / IK2_TILE_J;
        const int i_lo = (int)(a.grid.g0 / a.grid.ny), i_hi = (int)((a.grid.g1 - 1) / a.grid.ny);
        tile_i_begin = i_lo / IK2_TILE_I;
        blocks = (long long)(i_hi / IK2_TILE_I - tile_i_begin + 1) * tiles_j;
    } else {
        blocks = (a.grid.g1 - a.grid.g0 + IK2_THREADS - 1) / IK2_THREADS;
    }
    if (blocks <= 0) return cudaSuccess;
    const bool local = (a.obs_mask & (PGT_OBS_KINETIC | PGT_OBS_DENSITY)) != 0;
    if (local) k_ik_gather2<KERNEL, Q, true><<<(unsigned)blocks, IK2_THREADS, 0, st>>>(a, tile_i_begin);
    else k_ik_gather2<KERNEL, Q, false><<<(unsigned)blocks, IK2_THREADS, 0, st>>>(a, tile_i_begin);
    return cudaGetLastError();
}

template <int KERNEL>
static cudaError_t launch_ik2_q(const PgtIKArgs& a, cudaStream_t st)
{
    if (a.nq == 5) return launch_ik2_t<KERNEL, 5>(a, st);
    return launch_ik2_t<KERNEL, 0>(a, st);
}

static cudaError_t launch_ik2(const PgtIKArgs& a, cudaStream_t st)
{
    switch (a.cg.kernel) {
    case PGT_CG_OCTIC: return launch_ik2_q<PGT_CG_OCTIC>(a, st);
    case PGT_CG_RECT: return launch_ik2_q<PGT_CG_RECT>(a, st);
    default: return launch_ik2_q<PGT_CG_MOLLIFIER>(a, st);
    }
}

#include "pgt_ik_quad.inl"

cudaError_t pgt_launch_ik(const PgtIKArgs& a, int variant, cudaStream_t st, PgtLaunchCounter* lc)
{
    lc->n++;
    // automatic: the quad gather (unique bonds, records staged by bulk copies) wherever it applies -- Cartesian grids,
    // virial field only, a frame whose bond list carries primary masks; everything else (point lists, kinetic and
    // density fields) is the two-phase kernel
    if (variant != 2 && ik_quad_applicable(a)) {
        IKQCaps caps;
        size_t dyn = 0;
        if (ikq_pick_caps(a, &caps, &dyn)) return launch_ik_quad(a, caps, dyn, st);
    }
    return launch_ik2(a, st);
}

// ------------------------------------------------------------------------------------------
// Stage 4 with the adaptive bond quadrature (AdaptiveCGFunc, CoarseGrainFunction.h:94-134): one thread per grid point
// walks the reference's own loop (IrvingKirkwood.h:80-107) over the half-bond list and integrates
// Delta(|a + s b|^2) over s in [0, 1] with QAG / GK15 (pgt_qag15).  Not a throughput path: the interval store lives in
// local memory and the trip counts diverge; it exists so that cgfunc.*(mode="adaptive") runs on the device.
__global__ void __launch_bounds__(128) k_ik_adaptive(const PgtIKArgs A, int* __restrict__ eflags)
{
    const long long g = A.grid.g0 + (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= A.grid.g1) return;
    double gx, gy;
    if (A.grid.mode == 0) { gx = A.grid.xs[g / A.grid.ny]; gy = A.grid.ys[g % A.grid.ny]; }
    else { gx = A.grid.pts[3 * g]; gy = A.grid.pts[3 * g + 1]; }
    const PgtFrameView& fr = A.fr;
    const PgtCG cg = A.cg;
    const double R = cg.rcut, R2 = cg.rcutsq;
    const int ncx = A.cells.ncx, ncy = A.cells.ncy;
    const double padx = 1e-9 * (R + fabs(gx) + 1.0), pady = 1e-9 * (R + fabs(gy) + 1.0);
    const int ux0 = max(__double2int_rd((gx - R - padx - A.box.lo[0]) * A.cells.inv_cwx), -ncx);
    const int ux1 = min(__double2int_rd((gx + R + padx - A.box.lo[0]) * A.cells.inv_cwx), 2 * ncx - 1);
    const int uy0 = max(__double2int_rd((gy - R - pady - A.box.lo[1]) * A.cells.inv_cwy), -ncy);
    const int uy1 = min(__double2int_rd((gy + R + pady - A.box.lo[1]) * A.cells.inv_cwy), 2 * ncy - 1);
    Seg seg[3];
    const int nseg = split_segments(ux0, ux1, ncx, seg);
    double axx = 0.0, axy = 0.0, ayy = 0.0;
    int bad = 0;
    for (int uy = uy0; uy <= uy1; ++uy) {
        int ky = 0, row = uy;
        if (uy < 0) { ky = +1; row += ncy; } else if (uy >= ncy) { ky = -1; row -= ncy; }
        const double cpy = gy + ky * A.box.L[1];
        for (int s = 0; s < nseg; ++s) {
            const double cpx = gx + seg[s].k * A.box.L[0];
            const int qb = fr.cell_start[row * ncx + seg[s].c0], qe = fr.cell_start[row * ncx + seg[s].c1 + 1];
            for (int q = qb; q < qe; ++q) {
                const double ax = fr.sx[q] - cpx, ay = fr.sy[q] - cpy;
                if (ax * ax + ay * ay > R2) continue;
                for (int e = fr.hb_off[q]; e < fr.hb_off[q + 1]; ++e) {
                    const double2 b = fr.hb_b[e];
                    int st = 0;
                    const double B = pgt_qag15([&](double t) {
                        const double px = ax + t * b.x, py = ay + t * b.y;
                        return pgt_cg_delta_exact(cg, px * px + py * py);
                    }, 0.0, 1.0, A.qag_epsabs, A.qag_epsrel, A.qag_limit, &st);
                    if (st) bad |= (st == PGT_QAG_CAPACITY) ? PGT_EFLAG_QAG_CAPACITY : PGT_EFLAG_QAG_FAILED;
                    const double wf = 0.5 * B * fr.hb_cf[e].y;      // val += 0.5 * bondval * (f rij (x) rij)
                    const double tx = wf * b.x, ty = wf * b.y;
                    axx += tx * b.x;
                    axy += tx * b.y;
                    ayy += ty * b.y;
                }
            }
        }
    }
    if (bad) atomicOr(eflags, bad);
    *reinterpret_cast<double4*>(A.out_virial + 4 * (g - A.grid.g0)) = make_double4(axx, axy, axy, ayy);
}

cudaError_t pgt_launch_ik_adaptive(const PgtIKArgs& a, int* eflags, cudaStream_t st, PgtLaunchCounter* lc)
{
    const long long n = a.grid.g1 - a.grid.g0;
    if (n <= 0) return cudaSuccess;
    k_ik_adaptive<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(a, eflags);
    lc->n++;
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------
// ElasticVirialStress (ListofCommonObservables.h:86-123): val(a,b) += F1[a] w[b] - F[a] r[b] with r the pair separation the
// caller hands in, w = (disp_j - disp_i) + r, F = fdivr(|r|) r, F1 = fdivr(|w|) w.  Pairs are ALL neighbours within max_rcut
// (a pair beyond its own cut-off at r may carry force at w), so these kernels run the neighbour search themselves
// instead of reading the half-bond list.  Displacements are rows of the ORIGINAL particle order (setDisplacement).
template <int MODEL>
__device__ __forceinline__ void elastic_pair(const PgtPot& pot, double rx, double ry, double ux, double uy, double uz, double di,
                                             double dj, double& sxx, double& syx, double& sxy, double& syy)
{
    const double wx = ux + rx, wy = uy + ry;
    const double f = pgt_pair_eval_t<MODEL>(pot, rx * rx + ry * ry, di, dj, PGT_WANT_FDIVR);
    const double f1 = pgt_pair_eval_t<MODEL>(pot, wx * wx + wy * wy + uz * uz, di, dj, PGT_WANT_FDIVR);
    const double Fx = f * rx, Fy = f * ry, Gx = f1 * wx, Gy = f1 * wy;
    sxx = Gx * wx - Fx * rx;        // index a + 2 b
    syx = Gy * wx - Fy * rx;
    sxy = Gx * wy - Fx * ry;
    syy = Gy * wy - Fy * ry;
}

// GlobalElasticVirialStress: pairs id_i < id_j, r = +dx (ThermoCalculator.h:81-84); per-block partials of 4 sums
template <int MODEL>
__global__ void __launch_bounds__(THERMO_THREADS) k_thermo_elastic(PgtPot pot, PgtFrameView fr, PgtBox box, PgtCells cells, double rc2,
                                                                   const double* __restrict__ disp, double* __restrict__ partials)
{
    __shared__ double red[THERMO_THREADS / 32][4];
    double acc[4] = { 0.0, 0.0, 0.0, 0.0 };
    for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < fr.n; p += gridDim.x * blockDim.x) {
        const double xi = fr.sx[p], yi = fr.sy[p], di = fr.sd[p];
        const int oi = fr.sorig[p];
        const double uix = disp[3 * (size_t)oi], uiy = disp[3 * (size_t)oi + 1], uiz = disp[3 * (size_t)oi + 2];
        for_each_neighbor(fr, box, cells, p, xi, yi, rc2, [&](int q, double dx, double dy, double) {
            const int oj = __ldg(fr.sorig + q);
            if (!(oi < oj)) return;
            double sxx, syx, sxy, syy;
            elastic_pair<MODEL>(pot, dx, dy, disp[3 * (size_t)oj] - uix, disp[3 * (size_t)oj + 1] - uiy, disp[3 * (size_t)oj + 2] - uiz,
                                di, __ldg(fr.sd + q), sxx, syx, sxy, syy);
            acc[0] += sxx; acc[1] += syx; acc[2] += sxy; acc[3] += syy;
        });
    }
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const double s = warp_sum(acc[k]);
        if (lane == 0) red[w][k] = s;
    }
    __syncthreads();
    if (threadIdx.x < 4) {
        double s = 0.0;
        for (int ww = 0; ww < THERMO_THREADS / 32; ++ww) s += red[ww][threadIdx.x];
        partials[(size_t)blockIdx.x * 4 + threadIdx.x] = s;
    }
}

__global__ void __launch_bounds__(128) k_sum_partials4(const double* __restrict__ partials, int nblocks, double scale, double* __restrict__ out)
{
    const int k = threadIdx.x >> 5, lane = threadIdx.x & 31;
    double s = 0.0;
    for (int b = lane; b < nblocks; b += 32) s += partials[(size_t)b * 4 + k];
    s = warp_sum(s);
    if (lane == 0) out[k] = s * scale;                    // val * (1/vol)
}

void pgt_launch_thermo_elastic(const PgtPot& pot, PgtFrameView fr, PgtBox box, PgtCells cells, double max_rcut, const double* disp,
                               double inv_vol, double* partials, int blocks, double* out_dev, cudaStream_t st, PgtLaunchCounter* lc)
{
    const double rc2 = max_rcut * max_rcut;
    PGT_DISPATCH_MODEL(pot.model, (k_thermo_elastic<M><<<blocks, THERMO_THREADS, 0, st>>>(pot, fr, box, cells, rc2, disp, partials)));
    k_sum_partials4<<<1, 128, 0, st>>>(partials, blocks, inv_vol, out_dev);
    lc->n += 2;
}

// ElasticVirialStressField: the reference's loop (IrvingKirkwood.h:80-107), one thread per grid point.  The functor
// receives r = -dx there (:100), so w = (disp_j - disp_i) - dx: reproduced as written.  Fixed-point or adaptive rule.
template <int MODEL>
__global__ void __launch_bounds__(128) k_ik_elastic(const PgtPot pot, const PgtIKArgs A, const double* __restrict__ disp, double rc2,
                                                    int adaptive, double* __restrict__ out, int* __restrict__ eflags)
{
    const long long g = A.grid.g0 + (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= A.grid.g1) return;
    double gx, gy;
    if (A.grid.mode == 0) { gx = A.grid.xs[g / A.grid.ny]; gy = A.grid.ys[g % A.grid.ny]; }
    else { gx = A.grid.pts[3 * g]; gy = A.grid.pts[3 * g + 1]; }
    const PgtFrameView& fr = A.fr;
    const PgtCG cg = A.cg;
    const double R = cg.rcut, R2 = cg.rcutsq;
    const int ncx = A.cells.ncx, ncy = A.cells.ncy;
    const double padx = 1e-9 * (R + fabs(gx) + 1.0), pady = 1e-9 * (R + fabs(gy) + 1.0);
    const int ux0 = max(__double2int_rd((gx - R - padx - A.box.lo[0]) * A.cells.inv_cwx), -ncx);
    const int ux1 = min(__double2int_rd((gx + R + padx - A.box.lo[0]) * A.cells.inv_cwx), 2 * ncx - 1);
    const int uy0 = max(__double2int_rd((gy - R - pady - A.box.lo[1]) * A.cells.inv_cwy), -ncy);
    const int uy1 = min(__double2int_rd((gy + R + pady - A.box.lo[1]) * A.cells.inv_cwy), 2 * ncy - 1);
    Seg seg[3];
    const int nseg = split_segments(ux0, ux1, ncx, seg);
    double vxx = 0.0, vyx = 0.0, vxy = 0.0, vyy = 0.0;
    int bad = 0;
    for (int uy = uy0; uy <= uy1; ++uy) {
        int ky = 0, row = uy;
        if (uy < 0) { ky = +1; row += ncy; } else if (uy >= ncy) { ky = -1; row -= ncy; }
        const double cpy = gy + ky * A.box.L[1];
        for (int s = 0; s < nseg; ++s) {
            const double cpx = gx + seg[s].k * A.box.L[0];
            const int qb = fr.cell_start[row * ncx + seg[s].c0], qe = fr.cell_start[row * ncx + seg[s].c1 + 1];
            for (int p = qb; p < qe; ++p) {
                const double ax = fr.sx[p] - cpx, ay = fr.sy[p] - cpy;
                if (ax * ax + ay * ay > R2) continue;
                const double di = fr.sd[p];
                const int oi = fr.sorig[p];
                const double uix = disp[3 * (size_t)oi], uiy = disp[3 * (size_t)oi + 1], uiz = disp[3 * (size_t)oi + 2];
                for_each_neighbor(fr, A.box, A.cells, p, fr.sx[p], fr.sy[p], rc2, [&](int q, double dx, double dy, double) {
                    // bond function along r_i -> r_j (the reference reflects both vectors through x: same norms)
                    double B;
                    if (adaptive) {
                        int st = 0;
                        B = pgt_qag15([&](double t) {
                            const double px = ax + t * dx, py = ay + t * dy;
                            return pgt_cg_delta_exact(cg, px * px + py * py);
                        }, 0.0, 1.0, A.qag_epsabs, A.qag_epsrel, A.qag_limit, &st);
                        if (st) bad |= (st == PGT_QAG_CAPACITY) ? PGT_EFLAG_QAG_CAPACITY : PGT_EFLAG_QAG_FAILED;
                    } else {
                        B = 0.0;
                        for (int k = 0; k < A.nq; ++k) {
                            const double t = __ldg(A.gl_nodes + k), px = ax + t * dx, py = ay + t * dy;
                            B += __ldg(A.gl_weights + k) * pgt_cg_delta_exact(cg, px * px + py * py);
                        }
                    }
                    const int oj = __ldg(fr.sorig + q);
                    double sxx, syx, sxy, syy;
                    elastic_pair<MODEL>(pot, -dx, -dy, disp[3 * (size_t)oj] - uix, disp[3 * (size_t)oj + 1] - uiy,
                                        disp[3 * (size_t)oj + 2] - uiz, di, __ldg(fr.sd + q), sxx, syx, sxy, syy);
                    const double hb = 0.5 * B;                 // val += 0.5 * bondval * tmp
                    vxx += hb * sxx; vyx += hb * syx; vxy += hb * sxy; vyy += hb * syy;
                });
            }
        }
    }
    if (bad) atomicOr(eflags, bad);
    *reinterpret_cast<double4*>(out + 4 * (g - A.grid.g0)) = make_double4(vxx, vyx, vxy, vyy);
}

cudaError_t pgt_launch_ik_elastic(const PgtPot& pot, const PgtIKArgs& a, const double* disp, double max_rcut, int adaptive, double* out,
                                  int* eflags, cudaStream_t st, PgtLaunchCounter* lc)
{
    const long long n = a.grid.g1 - a.grid.g0;
    if (n <= 0) return cudaSuccess;
    const double rc2 = max_rcut * max_rcut;
    PGT_DISPATCH_MODEL(pot.model, (k_ik_elastic<M><<<(unsigned)((n + 127) / 128), 128, 0, st>>>(pot, a, disp, rc2, adaptive, out, eflags)));
    lc->n++;
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------
// Work counter for the roofline numerator (not on the hot path): one thread per grid point.
//   out[0] += U(x): unique force-carrying bonds with >= 1 endpoint in the disc (both tests exact)
//   out[1] += half-bond items the gather integrates,  out[2] += particles in the disc
__global__ void __launch_bounds__(128) k_ik_count(const PgtIKArgs A, unsigned long long* __restrict__ out)
{
    const long long g = A.grid.g0 + (long long)blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long u = 0, items = 0, parts = 0;
    if (g < A.grid.g1) {
        double gx, gy;
        if (A.grid.mode == 0) { gx = A.grid.xs[g / A.grid.ny]; gy = A.grid.ys[g % A.grid.ny]; }
        else { gx = A.grid.pts[3 * g]; gy = A.grid.pts[3 * g + 1]; }
        const PgtFrameView& fr = A.fr;
        const double R = A.cg.rcut, R2 = A.cg.rcutsq;
        const int ncx = A.cells.ncx, ncy = A.cells.ncy;
        const double padx = 1e-9 * (R + fabs(gx) + 1.0), pady = 1e-9 * (R + fabs(gy) + 1.0);
        int ux0 = max(__double2int_rd((gx - R - padx - A.box.lo[0]) * A.cells.inv_cwx), -ncx);
        int ux1 = min(__double2int_rd((gx + R + padx - A.box.lo[0]) * A.cells.inv_cwx), 2 * ncx - 1);
        int uy0 = max(__double2int_rd((gy - R - pady - A.box.lo[1]) * A.cells.inv_cwy), -ncy);
        int uy1 = min(__double2int_rd((gy + R + pady - A.box.lo[1]) * A.cells.inv_cwy), 2 * ncy - 1);
        Seg seg[3];
        const int nseg = split_segments(ux0, ux1, ncx, seg);
        for (int uy = uy0; uy <= uy1; ++uy) {
            int ky = 0, row = uy;
            if (uy < 0) { ky = +1; row += ncy; } else if (uy >= ncy) { ky = -1; row -= ncy; }
            const double cpy = gy + ky * A.box.L[1];
            for (int s = 0; s < nseg; ++s) {
                const double cpx = gx + seg[s].k * A.box.L[0];
                const int qb = fr.cell_start[row * ncx + seg[s].c0], qe = fr.cell_start[row * ncx + seg[s].c1 + 1];
                for (int q = qb; q < qe; ++q) {
                    const double ax = fr.sx[q] - cpx, ay = fr.sy[q] - cpy;
                    if (ax * ax + ay * ay > R2) continue;
                    ++parts;
                    for (int e = fr.hb_off[q]; e < fr.hb_off[q + 1]; ++e) {
                        ++items;
                        const int j = fr.hb_j[e];
                        const bool j_in = in_disc_exact(fr.sx[j], fr.sy[j], gx, gy, A.box, R2);
                        if (!j_in || fr.sorig[q] < fr.sorig[j]) ++u;
                    }
                }
            }
        }
    }
    // block reduce
    __shared__ unsigned long long red[3][4];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) {
        u += __shfl_xor_sync(FULL_MASK, u, d);
        items += __shfl_xor_sync(FULL_MASK, items, d);
        parts += __shfl_xor_sync(FULL_MASK, parts, d);
    }
    if (lane == 0) { red[0][w] = u; red[1][w] = items; red[2][w] = parts; }
    __syncthreads();
    if (threadIdx.x < 3) {
        unsigned long long s = red[threadIdx.x][0] + red[threadIdx.x][1] + red[threadIdx.x][2] + red[threadIdx.x][3];
        atomicAdd(out + threadIdx.x, s);     // integer counters: exact in any order
    }
}

void pgt_launch_ik_count(const PgtIKArgs& a, unsigned long long* out3, cudaStream_t st, PgtLaunchCounter* lc)
{
    const long long n = a.grid.g1 - a.grid.g0;
    if (n <= 0) return;
    k_ik_count<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(a, out3);
    lc->n++;
}

// ------------------------------------------------------------------------------------------
// Symmetric read-back: the 2x2 tensors of the stress fields are symmetric by construction (xy is stored twice), so the
// host needs 3 of the 4 doubles per grid point: packs (xx, xy, yy) for a 25 % smaller device-to-host copy.
__global__ void __launch_bounds__(256) k_pack_sym(const double4* __restrict__ in, double* __restrict__ out, long long n)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double4 v = in[i];
    out[3 * i] = v.x;
    out[3 * i + 1] = v.y;
    out[3 * i + 2] = v.w;
}

void pgt_launch_pack_sym(const double* in4, double* out3, long long n, cudaStream_t st, PgtLaunchCounter* lc)
{
    if (n <= 0) return;
    k_pack_sym<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(reinterpret_cast<const double4*>(in4), out3, n);
    lc->n++;
}

// ------------------------------------------------------------------------------------------
// Small results (counts, flags, the 25 thermo sums) reach the host through MAPPED page-locked memory written by
// this kernel, not through a device-to-host memcpy: a memcpy would queue in the copy engine behind a field
// read-back of another frame (134 MB at cfg5) and stall the stage that waits for it.
__global__ void k_publish(const void* __restrict__ s0, int n0, const void* __restrict__ s1, int n1, int* __restrict__ dst)
{
    const int t = threadIdx.x;
    if (t < n0) dst[t] = static_cast<const int*>(s0)[t];
    else if (t < n0 + n1) dst[t] = static_cast<const int*>(s1)[t - n0];
}

// A stage's count and error flags to the mapped slots dst[0], dst[1]; the flags -- and "the count exceeds the bound the host
// sized this frame's arrays for" -- are OR-ed into a sticky word that survives later frames (dst[4] mirrors it), so that a
// host that did not wait for this frame still learns of it at its next synchronisation point.
__global__ void k_publish_count(const int* __restrict__ count, const int* __restrict__ flags, long long bound, int* __restrict__ sticky,
                                int* __restrict__ dst, int* __restrict__ dst_sticky)
{
    if (threadIdx.x != 0) return;
    const int c = *count, f = *flags;
    dst[0] = c;
    dst[1] = f;
    const int s = *sticky | f | (((long long)c > bound || c < 0) ? PGT_EFLAG_BOUND : 0);
    *sticky = s;
    *dst_sticky = s;
}

void pgt_launch_publish_count(const int* count, const int* flags, long long bound, int* sticky, int* dst_mapped, int* dst_sticky,
                              cudaStream_t st, PgtLaunchCounter* lc)
{
    k_publish_count<<<1, 32, 0, st>>>(count, flags, bound, sticky, dst_mapped, dst_sticky);
    lc->n++;
}

void pgt_launch_publish(const void* s0, int nwords0, const void* s1, int nwords1, void* dst_mapped, cudaStream_t st, PgtLaunchCounter* lc)
{
    k_publish<<<1, 64, 0, st>>>(s0, nwords0, s1, nwords1, static_cast<int*>(dst_mapped));
    lc->n++;
}

// ------------------------------------------------------------------------------------------
// FP64 FMA peak probe: 8 independent dependent-chains per thread.
__global__ void __launch_bounds__(256) k_fp64_probe(double* out, int iters)
{
    double a0 = threadIdx.x * 1e-3, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
    const double m = 0.999999, c = 1e-7;
    for (int i = 0; i < iters; ++i) {
        a0 = fma(a0, m, c); a1 = fma(a1, m, c); a2 = fma(a2, m, c); a3 = fma(a3, m, c);
        a4 = fma(a4, m, c); a5 = fma(a5, m, c); a6 = fma(a6, m, c); a7 = fma(a7, m, c);
    }
    out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
}

void pgt_launch_fp64_probe(double* out, int iters, int blocks, int threads, cudaStream_t st, PgtLaunchCounter* lc)
{
    k_fp64_probe<<<blocks, threads, 0, st>>>(out, iters);
    lc->n++;
}
