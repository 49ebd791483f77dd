// pgt_ik_quad.inl -- stage 4, Cartesian grids: the "quad" gather (included by pgt_kernels.cu).
//
// Replaces IrvingKirkwood::compute (irvingkirkwood/IrvingKirkwood.h:74-110) for the virial field on Cartesian grids.
// A CTA owns a tile of 8 wx by 4 wy grid points (one thread per point for the discovery phases); a warp covers a compact
// 8 x 4 patch.  What bounds this kernel is (a) the FP64 pipe in the integration loop, (b) issue slots in discovery and
// (c) the shared-memory crossbar -- NOT the L1/texture path: a scattered per-lane global load costs ~2 L1 wavefront
// cycles per lane on sm_100, which is what held every formulation that read the records through L1 at ~1.8 ms (cfg5).
//
//  stage     The cells within R of the tile form a rectangle of cell rows; the frame is sorted by cell (row-major), so a
//            row of the rectangle is ONE contiguous run of particles and ONE contiguous run of half-bond records (two
//            when the tile straddles the periodic seam).  One warp issues a bulk asynchronous copy (cp.async.bulk, the
//            1-D TMA path: UBLKCP in SASS) per run and array -- positions (x, y), b = r_j - r_i, (|b|^2/R^2, f) -- that
//            complete on an mbarrier; meanwhile all warps build the small index tables (local cell starts, each
//            particle's record range and primary mask, each record's owner) from cell_start / hb_off.
//  phase 1   thread <-> grid point: the particles of MY cell window are tested exactly as Aboria does
//            (dx = r_p - (x + kL), dx^2 + dy^2 <= R^2, Search.h:186-193, 438-447); one 32-bit hit mask per window row.
//  phase 2   thread <-> grid point, "walk": one step = the next particle i of my disc (if the current one is used up),
//            then four of its half-bonds i -> j.  The reference visits a bond from each end that lies in the disc,
//            each time with weight 1/2 and the same line integral; here a bond is kept ONCE: always from its primary
//            end (sorted index i < j: bit of the particle's primary mask), from the other end only when the primary end
//            is outside the disc.  Whether j is inside is decided arithmetically, |a + b|^2 against R^2 with a guard
//            band; inside the band (rounding could matter) the exact Aboria test on j's stored position decides, so
//            the discrete weights (1, 1/2, 0) are the reference's bit for bit.  Survivors go to my list in shared
//            memory (2 bytes: record + both-ends-inside bit).  U(x) quadratures instead of ~1.5 U(x).
//  phase 3   FOUR LANES <-> grid point: the lanes of a quad stride over the point's list, so the eight lanes of a
//            128-bit shared-memory phase read (nearly) consecutive records instead of eight random ones (2.3-way bank
//            conflicts otherwise), and the work per lane is balanced to within 4 items.  a = r_i - x,
//            t(s) = c0 + s (c1 + s c2), Gauss-Legendre sum of the shape in GSL's order (FixedPointQuadrature.h:35-45),
//            B f b(x)b into three accumulators; a two-step shuffle reduction hands the quad's sum to the point's thread.
//            Phases 2 and 3 alternate on warp-uniform conditions until every list of the warp is drained.
//
// Determinism: every in-grid point of a live warp is computed whatever the shard, list capacities come from frame-wide
// statistics only, and a point's items are summed in a fixed order -> the field is bit-identical for every sharding.
// A tile that does not meet the limits runs `ik_slow_path` (plain half-bond formulation from global memory).
#define IKQ_PX 8
#define IKQ_PY 4
#define IKQ_MAXROWS 32
#define IKQ_MAXRUNS (2 * IKQ_MAXROWS)
#define IKQ_MAX_THREADS 256               // 2 CTAs of 256 threads per SM (shared memory bound): up to 128 registers each
#define IKQ_SEC 0x8000u                   // list entry: kept from the bond's secondary end (weight 1/2, no far-end test)
#define IKQ_MAX_NREC 0x1FFF               // a list entry / a particle's record range addresses 13 bits
#define IKQ_STEP_ROOM 6                   // list slots one walk step may touch (4 primary + 2 secondary)

struct IKQCaps {
    int wx, wy, wy_log2;                 // warps per tile along i (x) and j (y; a power of two)
    int rows_cap;                        // cell rows of the rectangle
    int lcs_stride;                      // entries per row of the local cell-start table (columns of the rectangle + 1)
    int nloc_cap;                        // particles of the rectangle
    int nrec_cap;                        // half-bond records of the rectangle
    int buf_cap;                         // entries of a thread's buffer: disc particles from the top down, list items from the bottom up
    int buf_stride;                      // 16-bit words between two threads' buffers (2 x odd: conflict-free columns)
    int rank_bits;                       // bits of a list length
    double r2lo, r2hi;                   // |a + b|^2 <= r2lo: inside; > r2hi: outside; between: exact test
    double r2keep;                       // phase 2 keeps a secondary item when |a + b|^2 > R^2 - band / 2
    double t1lo, t1hi;                   // the same band for t(1) = |a + b|^2 / R^2
    float r2lo32, r2hi32;                // the band of the single-precision disc test of phase 1 (tile-relative coordinates)
};

struct IKQStatic {
    int ps[IKQ_MAXRUNS];                 // first sorted particle of the run
    int np[IKQ_MAXRUNS];                 // particles in the run
    int h0[IKQ_MAXRUNS];                 // first half-bond record of the run (global)
    int lp_base[IKQ_MAXRUNS + 1];        // tile-local index of the run's first particle
    int lr_base[IKQ_MAXRUNS + 1];        // tile-local index of the run's first record
    double2 shift[IKQ_MAXRUNS];          // (kx Lx, ky Ly) of the run's periodic image
    int bad;                             // a particle of the tile has records the packed range cannot describe
};

// ---- mbarrier / bulk-copy primitives (PTX ISA 8.x; sm_90+)
__device__ __forceinline__ unsigned ikq_smem(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void ikq_mbar_init(unsigned bar, unsigned count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void ikq_mbar_expect_tx(unsigned bar, unsigned bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void ikq_bulk_g2s(unsigned dst, const void* src, unsigned bytes, unsigned bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src),
                 "r"(bytes), "r"(bar)
                 : "memory");
}
__device__ __forceinline__ void ikq_mbar_wait(unsigned bar, unsigned parity)
{
    asm volatile("{\n"
                 ".reg .pred P1;\n"
                 "IKQ_WAIT:\n"
                 "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
                 "@P1 bra IKQ_DONE;\n"
                 "bra IKQ_WAIT;\n"
                 "IKQ_DONE:\n"
                 "}" ::"r"(bar), "r"(parity)
                 : "memory");
}

// The reference's own formulation for one grid point, from global memory (no tile limits): every half-bond of every
// particle in the disc, weight 1/2 (IrvingKirkwood.h:80-107).
template <int KERNEL, int Q>
__device__ __noinline__ void ik_slow_path(const PgtIKArgs& A, double gx, double gy, double& oxx, double& oxy, double& oyy)
{
    const PgtFrameView& fr = A.fr;
    const double R = A.cg.rcut, R2 = A.cg.rcutsq, iR2 = A.inv_rcutsq;
    const int ncx = A.cells.ncx, ncy = A.cells.ncy;
    const double padx = 1e-9 * (R + fabs(gx) + 1.0), pady = 1e-9 * (R + fabs(gy) + 1.0);
    const int ux0 = max(__double2int_rd((gx - R - padx - A.box.lo[0]) * A.cells.inv_cwx), -ncx);
    const int ux1 = min(__double2int_rd((gx + R + padx - A.box.lo[0]) * A.cells.inv_cwx), 2 * ncx - 1);
    const int uy0 = max(__double2int_rd((gy - R - pady - A.box.lo[1]) * A.cells.inv_cwy), -ncy);
    const int uy1 = min(__double2int_rd((gy + R + pady - A.box.lo[1]) * A.cells.inv_cwy), 2 * ncy - 1);
    Seg seg[3];
    const int nseg = split_segments(ux0, ux1, ncx, seg);
    double axx = 0.0, axy = 0.0, ayy = 0.0;
    for (int uy = uy0; uy <= uy1; ++uy) {
        int ky = 0, row = uy;
        if (uy < 0) { ky = +1; row += ncy; } else if (uy >= ncy) { ky = -1; row -= ncy; }
        const double cpy = gy + ky * A.box.L[1];
        for (int s = 0; s < nseg; ++s) {
            const double cpx = gx + seg[s].k * A.box.L[0];
            const int qb = fr.cell_start[row * ncx + seg[s].c0], qe = fr.cell_start[row * ncx + seg[s].c1 + 1];
            for (int q = qb; q < qe; ++q) {
                const double ax = fr.sx[q] - cpx, ay = fr.sy[q] - cpy;
                const double d2 = ax * ax + ay * ay;
                if (d2 > R2) continue;
                const double a2x = ax * (2.0 * iR2), a2y = ay * (2.0 * iR2), c0 = d2 * iR2;
                for (int e = fr.hb_off[q]; e < fr.hb_off[q + 1]; ++e) {
                    const double2 b = fr.hb_b[e], cf = fr.hb_cf[e];
                    const double c1 = fma(a2x, b.x, a2y * b.y);
                    const double wf = bond_quadrature<KERNEL, Q>(A, c0, c1, cf.x) * cf.y * 0.5;
                    const double tx = wf * b.x, ty = wf * b.y;
                    axx = fma(tx, b.x, axx);
                    axy = fma(tx, b.y, axy);
                    ayy = fma(ty, b.y, ayy);
                }
            }
        }
    }
    oxx = axx; oxy = axy; oyy = ayy;
}

// The run of the rectangle that holds tile-local particle l.
__device__ __forceinline__ int ikq_run_of(const IKQStatic& S, int l, int nruns)
{
    int run = 0;
    while (run + 1 < nruns && l >= S.lp_base[run + 1]) ++run;
    return run;
}

// Are the far ends of the records k + u (u in `which`; particle l of the rectangle) found by the reference's search
// around (gx, gy)?  The exact Aboria test on j's stored position; only reached inside the guard band of the arithmetic
// test.  Returns `inm` with the bits of `which` replaced by the exact answers.
__device__ __noinline__ unsigned ikq_exact_far_ends(const PgtIKArgs& A, const IKQStatic& S, int l, int k, unsigned which, unsigned inm,
                                                    double gx, double gy, int nruns)
{
    const int run = ikq_run_of(S, l, nruns);                       // the run that holds particle l (and so record k)
    for (int u = 0; u < 4; ++u) {
        if (!((which >> u) & 1u)) continue;
        const int e = S.h0[run] + (k + u - S.lr_base[run]);
        const int j = __ldg(A.fr.hb_j + e);
        const bool in = in_disc_exact(__ldg(A.fr.sx + j), __ldg(A.fr.sy + j), gx, gy, A.box, A.cg.rcutsq);
        inm = (inm & ~(1u << u)) | ((in ? 1u : 0u) << u);
    }
    return inm;
}

// Is tile-local particle l found by the reference's search around (gx, gy)?  Exact test on its stored position; only
// reached inside the guard band of phase 1's single-precision test.
__device__ __noinline__ bool ikq_exact_hit(const PgtIKArgs& A, const IKQStatic& S, int l, double gx, double gy, int nruns)
{
    const int run = ikq_run_of(S, l, nruns);
    const int p = S.ps[run] + (l - S.lp_base[run]);
    return in_disc_exact(__ldg(A.fr.sx + p), __ldg(A.fr.sy + p), gx, gy, A.box, A.cg.rcutsq);
}

// shared-memory accesses by 32-bit address (the walk keeps one base register per table)
__device__ __forceinline__ unsigned ikq_lds_u16(unsigned addr) { unsigned short v; asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(addr)); return v; }
__device__ __forceinline__ unsigned ikq_lds_u32(unsigned addr) { unsigned v; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr)); return v; }
template <int OFF>
__device__ __forceinline__ void ikq_sts_u16(unsigned addr, unsigned v) { asm volatile("st.shared.u16 [%0+%2], %1;" ::"r"(addr), "h"((unsigned short)v), "n"(OFF) : "memory"); }
template <int OFF>
__device__ __forceinline__ double2 ikq_lds_f64x2(unsigned addr)
{
    double2 v;
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2+%3];" : "=d"(v.x), "=d"(v.y) : "r"(addr), "n"(OFF));
    return v;
}
__device__ __forceinline__ float2 ikq_lds_f32x2(unsigned addr)
{
    float2 v;
    asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(addr));
    return v;
}

// One quad's share of a grid point's list with the far-end tests of the primary items done exactly (the reference's
// search on j's stored position): the rare redo of a phase-3 pass in which some t(1) fell inside the guard band.
template <int KERNEL, int Q>
__device__ __noinline__ void ikq_pass_exact(const PgtIKArgs& A, const IKQStatic& S, const IKQCaps& caps, unsigned lp_addr, int n_p, int sub,
                                            double gxp, double gyp, int nruns, double& oxx, double& oxy, double& oyy)
{
    extern __shared__ __align__(16) unsigned char ikq_dyn[];
    const double2* const s_b = reinterpret_cast<const double2*>(ikq_dyn);
    const double2* const s_cf = s_b + caps.nrec_cap;
    const double2* const s_pos = s_cf + caps.nrec_cap;
    const unsigned short* const s_own = reinterpret_cast<const unsigned short*>(ikq_dyn + (size_t)caps.nrec_cap * 32 + (size_t)caps.nloc_cap * 28) +
                                        (size_t)blockDim.x * caps.buf_stride;
    const double iR2 = A.inv_rcutsq, iR2x2 = 2.0 * A.inv_rcutsq;
    double axx = 0.0, axy = 0.0, ayy = 0.0;
    for (int s = sub; s < n_p; s += 4) {
        const unsigned it = ikq_lds_u16(lp_addr + 2u * (unsigned)s);
        const unsigned kk = it & (unsigned)IKQ_MAX_NREC;
        const int l = s_own[kk];
        const double2 p = s_pos[l], b = s_b[kk], cf = s_cf[kk];
        const double pax = p.x - gxp, pay = p.y - gyp;
        const double c0 = fma(pax, pax, pay * pay) * iR2;
        const double c1 = fma(pax, b.x, pay * b.y) * iR2x2;
        const double t1 = (c0 + cf.x) + c1;
        // far end clearly outside: one visit (1/2).  Otherwise a secondary item is always decided exactly (far end inside:
        // the bond is integrated from its primary end, not here), a primary item inside the band too
        const bool sec = (it & IKQ_SEC) != 0u;
        double wh = 0.5;
        if (!(t1 > caps.t1hi)) {
            if (sec || t1 > caps.t1lo) {
                const bool in = (ikq_exact_far_ends(A, S, l, (int)kk, 1u, 0u, gxp, gyp, nruns) & 1u) != 0u;
                wh = sec ? (in ? 0.0 : 0.5) : (in ? 1.0 : 0.5);
            } else {
                wh = 1.0;
            }
        }
        const double wf = ik_quadrature<KERNEL, Q>(A, c0, c1, cf.x) * cf.y;
        const double tx = (wf * wh) * b.x, ty = (wf * wh) * b.y;
        axx = fma(tx, b.x, axx);
        axy = fma(tx, b.y, axy);
        ayy = fma(ty, b.y, ayy);
    }
    oxx = axx; oxy = axy; oyy = ayy;
}

template <int KERNEL, int Q>
__global__ void __launch_bounds__(IKQ_MAX_THREADS, 2) k_ik_quad(const __grid_constant__ PgtIKArgs A, const int tile_i_begin, const IKQCaps caps)
{
    __shared__ IKQStatic S;
    __shared__ __align__(8) unsigned long long s_bar;
    __shared__ unsigned char s_perm[IKQ_MAX_THREADS];      // per warp: lane of the point whose list has a given rank
    extern __shared__ __align__(16) unsigned char ikq_dyn[];
    const int T = blockDim.x;
    double2* const s_b = reinterpret_cast<double2*>(ikq_dyn);
    double2* const s_cf = s_b + caps.nrec_cap;
    double2* const s_pos = s_cf + caps.nrec_cap;
    float2* const s_p32 = reinterpret_cast<float2*>(s_pos + caps.nloc_cap);             // tile-relative, single precision (phase 1)
    unsigned* const s_rng = reinterpret_cast<unsigned*>(s_p32 + caps.nloc_cap);          // first record | secondaries << 13 | primaries << 17
    unsigned short* const s_buf = reinterpret_cast<unsigned short*>(s_rng + caps.nloc_cap);    // [thread][slot]: disc particles / bonds to integrate
    unsigned short* const s_own = s_buf + T * caps.buf_stride;                           // local index of i per record
    unsigned short* const s_lcs = s_own + caps.nrec_cap;                                 // [row][lcs_stride]: first particle of a cell

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const PgtFrameView& fr = A.fr;
    const double R = A.cg.rcut, iR2 = A.inv_rcutsq;
    const int ncx = A.cells.ncx, ncy = A.cells.ncy;
    const unsigned bar = ikq_smem(&s_bar);
    const int LCS = caps.lcs_stride;

    // ---- tile and my grid point: a warp covers a compact 8 x 4 patch
    const int TI = IKQ_PX * caps.wx, TJ = IKQ_PY * caps.wy;
    const int i0 = (tile_i_begin + (int)blockIdx.y) * TI, j0 = (int)blockIdx.x * TJ;
    const int i1 = min(i0 + TI, A.grid.nx) - 1, j1 = min(j0 + TJ, A.grid.ny) - 1;
    const int gi = i0 + IKQ_PX * (warp >> caps.wy_log2) + (lane >> 2), gj = j0 + IKQ_PY * (warp & (caps.wy - 1)) + (lane & 3);   // wy is 1, 2 or 4
    // every in-grid point of a warp that stores anything is computed, whatever the shard: a point's summation order
    // (and so its bits) must not depend on how the grid is split over GPUs
    const bool ingrid = gi <= i1 && gj <= j1;
    bool store = false;
    long long g = 0;
    double gx = 0.0, gy = 0.0;
    if (ingrid) {
        g = (long long)gi * A.grid.ny + gj;
        store = g >= A.grid.g0 && g < A.grid.g1;
        gx = __ldg(A.grid.xs + gi);
        gy = __ldg(A.grid.ys + gj);
    }
    const bool warp_live = __any_sync(FULL_MASK, store);
    if (tid == 0) { ikq_mbar_init(bar, 1); S.bad = 0; }
    if (!__syncthreads_or(store ? 1 : 0)) return;           // no point of this tile lies in the shard

    // ---- cell rectangle of the tile (unwrapped cell coordinates), same padding rule as a single point's window
    double xmin = __ldg(A.grid.xs + i0), xmax = xmin, ymin = __ldg(A.grid.ys + j0), ymax = ymin;
    if (A.grid.monotone) {                                 // non-decreasing axes (np.linspace): the end points bound the tile
        xmax = __ldg(A.grid.xs + i1);
        ymax = __ldg(A.grid.ys + j1);
    } else {
        for (int i = i0 + 1; i <= i1; ++i) { const double v = __ldg(A.grid.xs + i); xmin = fmin(xmin, v); xmax = fmax(xmax, v); }
        for (int j = j0 + 1; j <= j1; ++j) { const double v = __ldg(A.grid.ys + j); ymin = fmin(ymin, v); ymax = fmax(ymax, v); }
    }
    const double tpx = 1e-9 * (R + fmax(fabs(xmin), fabs(xmax)) + 1.0), tpy = 1e-9 * (R + fmax(fabs(ymin), fabs(ymax)) + 1.0);
    const int ux0 = __double2int_rd((xmin - R - tpx - A.box.lo[0]) * A.cells.inv_cwx);
    const int ux1 = __double2int_rd((xmax + R + tpx - A.box.lo[0]) * A.cells.inv_cwx);
    const int uy0 = __double2int_rd((ymin - R - tpy - A.box.lo[1]) * A.cells.inv_cwy);
    const int uy1 = __double2int_rd((ymax + R + tpy - A.box.lo[1]) * A.cells.inv_cwy);
    const int ncols = ux1 - ux0 + 1, nrows = uy1 - uy0 + 1;
    // limits; the rectangle plus a bond's reach must not wrap onto itself (one image of every particle)
    bool fast = ncols + 1 <= LCS && nrows <= caps.rows_cap && ncols + 2 * A.cells.reach + 1 <= ncx &&
                nrows + 2 * A.cells.reach + 1 <= ncy && ux0 >= -ncx && ux1 < 2 * ncx && uy0 >= -ncy && uy1 < 2 * ncy;

    Seg seg[3];
    int nseg = 0;
    if (fast) nseg = split_segments(ux0, ux1, ncx, seg);
    if (nseg > 2) fast = false;
    const int nruns = 2 * nrows;
    const int csplit = (nseg == 2) ? (seg[0].c1 - seg[0].c0 + 1) : ncols;   // column at which the second segment starts

    int nloc = 0, nrec_total = 0;
    if (fast) {
        // ---- runs: (row, segment) -> contiguous particles [ps, ps+np) and records
        for (int run = tid; run < nruns; run += T) {
            const int r = run >> 1, s = run & 1;
            int uy = uy0 + r, ky = 0;
            if (uy < 0) { ky = +1; uy += ncy; } else if (uy >= ncy) { ky = -1; uy -= ncy; }
            int ps = 0, pe = 0, h0 = 0, h1 = 0, kx = 0;
            if (s < nseg) {
                ps = __ldg(fr.cell_start + uy * ncx + seg[s].c0);
                pe = __ldg(fr.cell_start + uy * ncx + seg[s].c1 + 1);
                kx = seg[s].k;
                if (pe > ps) { h0 = __ldg(fr.hb_off + ps); h1 = __ldg(fr.hb_off + pe); }
            }
            S.ps[run] = ps;
            S.np[run] = pe - ps;
            S.h0[run] = h0;
            S.lr_base[run] = h1 - h0;                      // counts for now, scanned below
            S.shift[run] = make_double2(kx * A.box.L[0], ky * A.box.L[1]);
        }
        __syncthreads();
        if (warp == 0) {                                   // exclusive scans over <= 64 runs: lane <-> row
            int cp0 = 0, cp1 = 0, cr0 = 0, cr1 = 0;
            if (lane < nrows) { cp0 = S.np[2 * lane]; cp1 = S.np[2 * lane + 1]; cr0 = S.lr_base[2 * lane]; cr1 = S.lr_base[2 * lane + 1]; }
            int sp = cp0 + cp1, sr = cr0 + cr1;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int tp = __shfl_up_sync(FULL_MASK, sp, d), tr = __shfl_up_sync(FULL_MASK, sr, d);
                if (lane >= d) { sp += tp; sr += tr; }
            }
            if (lane < nrows) {
                S.lp_base[2 * lane] = sp - cp0 - cp1;
                S.lp_base[2 * lane + 1] = sp - cp1;
                S.lr_base[2 * lane] = sr - cr0 - cr1;
                S.lr_base[2 * lane + 1] = sr - cr1;
            }
            if (lane == nrows - 1) { S.lp_base[nruns] = sp; S.lr_base[nruns] = sr; }
        }
        __syncthreads();
        nloc = S.lp_base[nruns];
        nrec_total = S.lr_base[nruns];
        // uniform: every thread reads the same totals; two spare records: the walk loads a pair of record slots
        if (nloc > caps.nloc_cap || nrec_total + 2 > caps.nrec_cap) fast = false;
    }

    double txx = 0.0, txy = 0.0, tyy = 0.0;
    const double out_scale = A.cg.norm;                    // weights 1 / 0.5 already carry VirialStress::compute_cg's 1/2
    double* const out_ptr = A.out_virial + 4 * (g - A.grid.g0);

    if (fast) {
        // ---- one warp starts the bulk copies: a run's positions and its records are contiguous in global memory
        if (warp == 0) {
            if (lane == 0) ikq_mbar_expect_tx(bar, 16u * (unsigned)(nloc + 2 * nrec_total));
            __syncwarp();
            for (int run = lane; run < nruns; run += 32) {
                const int np = S.np[run], lrb = S.lr_base[run], nr = S.lr_base[run + 1] - lrb, h0 = S.h0[run];
                if (np > 0) ikq_bulk_g2s(ikq_smem(s_pos + S.lp_base[run]), fr.sxy + S.ps[run], 16u * (unsigned)np, bar);
                if (nr > 0) {
                    ikq_bulk_g2s(ikq_smem(s_b + lrb), fr.hb_b + h0, 16u * (unsigned)nr, bar);
                    ikq_bulk_g2s(ikq_smem(s_cf + lrb), fr.hb_cf + h0, 16u * (unsigned)nr, bar);
                }
            }
        }
        // ---- meanwhile: the index tables.  Local cell starts
        for (int r = warp; r < nrows; r += (T >> 5)) {
            int uy = uy0 + r;
            if (uy < 0) uy += ncy; else if (uy >= ncy) uy -= ncy;
            for (int c = lane; c <= ncols; c += 32) {
                int v;
                if (c == ncols) v = S.lp_base[2 * r + 2];
                else {
                    const int s = c >= csplit;
                    const int cell = uy * ncx + (s ? seg[1].c0 + (c - csplit) : seg[0].c0 + c);
                    v = S.lp_base[2 * r + s] + (__ldg(fr.cell_start + cell) - S.ps[2 * r + s]);
                }
                s_lcs[r * LCS + c] = (unsigned short)v;
            }
        }
        // a warp takes a run: each particle's record range (secondary records first, then primary ones), each record's owner
        const int nwarps = T >> 5;
        int bad = 0;
        for (int run = warp; run < nruns; run += nwarps) {
            const int ps = S.ps[run], np = S.np[run], lpb = S.lp_base[run], h0 = S.h0[run], lrb = S.lr_base[run];
            for (int o = lane; o < np; o += 32) {
                const int p = ps + o, l = lpb + o;
                const int k0 = lrb + (__ldg(fr.hb_off + p) - h0), k1 = lrb + (__ldg(fr.hb_off + p + 1) - h0);
                const unsigned pm = __ldg(fr.hb_pm + p);
                const int n = k1 - k0, npr = __popc(pm), nsec = n - npr;
                // the primary records must be the last npr of the n (k_bonds_find's visiting order guarantees it)
                if (n > 15 || nsec < 0 || pm != (((1u << n) - 1u) & ~((1u << nsec) - 1u))) bad = 1;
                s_rng[l] = (unsigned)k0 | ((unsigned)(nsec & 15) << 13) | ((unsigned)(npr & 15) << 17);
                for (int k = k0; k < k1; ++k) s_own[k] = (unsigned short)l;
            }
        }
        if (bad) S.bad = 1;
        __syncthreads();
        ikq_mbar_wait(bar, 0);                             // the copies have landed (and nothing is in flight when we leave)
        if (S.bad) fast = false;
    }

    if (!fast) {
        if (store) {
            ik_slow_path<KERNEL, Q>(A, gx, gy, txx, txy, tyy);
            *reinterpret_cast<double4*>(out_ptr) = make_double4(out_scale * txx, out_scale * txy, out_scale * txy, out_scale * tyy);
        }
        return;
    }

    // ---- positions of the tile's periodic image, r - kL (used for the quadrature and the guarded arithmetic tests only,
    //      never for an exact membership test), and their single-precision copy relative to the tile's corner
    for (int run = warp; run < nruns; run += (T >> 5)) {
        const int np = S.np[run], lpb = S.lp_base[run];
        const double2 sh = S.shift[run];
        for (int o = lane; o < np; o += 32) {
            double2 p = s_pos[lpb + o];
            p.x -= sh.x; p.y -= sh.y;
            s_pos[lpb + o] = p;
            s_p32[lpb + o] = make_float2((float)(p.x - xmin), (float)(p.y - ymin));
        }
    }
    __syncthreads();

    // ================= phase 1: the particles of my disc, as a list of tile-local indices =================
    // One flat loop over the candidates of all my window rows (a row's candidates are consecutive local indices), so
    // the 32 searches of a warp cost max-over-lanes of the TOTAL candidate count, not the sum of per-row maxima.
    // The test runs in single precision on tile-relative coordinates (the FP64 pipe is the integration loop's); a
    // candidate inside the guard band of that test is decided by the exact Aboria test on its stored position.
    const bool active = ingrid && warp_live;
    bool slow = false;                                     // my disc does not fit the buffer: I take the slow path
    const unsigned buf_addr = ikq_smem(s_buf + tid * caps.buf_stride);     // items grow upwards from here
    const unsigned buf_top = buf_addr + 2u * (unsigned)caps.buf_cap;       // disc particles grow downwards from here
    int nh = 0;                                            // hits
    if (active) {
        const double padx = 1e-9 * (R + fabs(gx) + 1.0), pady = 1e-9 * (R + fabs(gy) + 1.0);
        const int cx0 = max(__double2int_rd((gx - R - padx - A.box.lo[0]) * A.cells.inv_cwx), ux0) - ux0;
        const int cx1 = min(__double2int_rd((gx + R + padx - A.box.lo[0]) * A.cells.inv_cwx), ux1) - ux0;
        const int ry0 = max(__double2int_rd((gy - R - pady - A.box.lo[1]) * A.cells.inv_cwy), uy0) - uy0;
        const int ry1 = min(__double2int_rd((gy + R + pady - A.box.lo[1]) * A.cells.inv_cwy), uy1) - uy0;
        const float g32x = (float)(gx - xmin), g32y = (float)(gy - ymin);
        const float r2lo32 = caps.r2lo32, r2hi32 = caps.r2hi32;
        const unsigned p32_addr = ikq_smem(s_p32);
        const int hit_max = caps.buf_cap - IKQ_STEP_ROOM - 2;
        int r = ry0 - 1, l = 0, lend = 0;
        unsigned hp = buf_top;
        const unsigned hp_min = buf_top - 2u * (unsigned)(hit_max + 1);   // reaching it means the disc does not fit
        unsigned nu = 0u;
        for (;;) {
            if (l >= lend) {                               // next window row (empty rows cost one turn)
                if (++r > ry1) break;
                l = s_lcs[r * LCS + cx0];
                lend = s_lcs[r * LCS + cx1 + 1];
                continue;
            }
            const float2 p = ikq_lds_f32x2(p32_addr + 8u * (unsigned)l);
            const float dx = p.x - g32x, dy = p.y - g32y;
            const float d2 = fmaf(dx, dx, dy * dy);
            // straight-line: the slot below the list's end is written always, the end moves for candidates only; one
            // inside the guard band is marked (bit 15) and decided after the loop
            const unsigned mark = d2 > r2lo32 ? 0x8000u : 0u;
            ikq_sts_u16<-2>(hp, (unsigned)l | mark);
            if (d2 <= r2hi32) { hp = max(hp - 2u, hp_min); nu |= mark; }
            ++l;
        }
        if (nu) {                                          // rare: exact tests, the list is compacted in place
            unsigned rp = buf_top, wq = buf_top;
            while (rp > hp) {
                rp -= 2u;
                const unsigned e = ikq_lds_u16(rp);
                if (!(e & 0x8000u) || ikq_exact_hit(A, S, (int)(e & 0x7FFFu), gx, gy, nruns)) { wq -= 2u; ikq_sts_u16<0>(wq, e & 0x7FFFu); }
            }
            hp = wq;
        }
        nh = (int)((buf_top - hp) >> 1);
        if (nh > hit_max) { slow = true; nh = 0; }
    }
    if (!warp_live) return;

    // ================= phases 2 and 3, alternating; warp-uniform control only =================
    // explicit 32-bit shared addresses in the hot loops: one register per table, immediate offsets for the slots
    const unsigned b_addr = ikq_smem(s_b), cf_addr = ikq_smem(s_cf), pos_addr = ikq_smem(s_pos), rng_addr = ikq_smem(s_rng);
    const unsigned own_addr = ikq_smem(s_own), buf_base = ikq_smem(s_buf);
    const double r2lo = caps.r2keep, t1lo = caps.t1lo, t1hi = caps.t1hi;
    unsigned hp = buf_top - 2u * (unsigned)nh;             // next disc particle to walk (the last one found comes first)
    int ks = 0, ns_rem = 0, kp = 0, np_rem = 0, lcur = 0;
    double ax = 0.0, ay = 0.0;
    const double iR2x2 = 2.0 * iR2;
    const int sub = lane & 3;

    for (;;) {
        // ---- phase 2: walk.  One step = (if my particle is used up) the next particle i of my disc, then up to four of
        //      its primary records -- always items: the bond is integrated once, from this end -- and up to two of its
        //      secondary records, items only when the bond's primary end lies outside the disc (|a + b|^2 against R^2
        //      with a guard band).  Straight-line: slots are written unconditionally and the write pointer moves past
        //      the kept ones.  A vote per step: stop when some buffer is nearly full or every lane has finished
        unsigned wp = buf_addr;                                // my list's write pointer
        int state;
        do {
            if (ns_rem <= 0 && np_rem <= 0 && hp < buf_top) {
                lcur = (int)ikq_lds_u16(hp);
                hp += 2u;
                const unsigned rng = ikq_lds_u32(rng_addr + 4u * (unsigned)lcur);
                const double2 p = ikq_lds_f64x2<0>(pos_addr + 16u * (unsigned)lcur);
                ks = (int)(rng & 0x1FFFu);
                ns_rem = (int)((rng >> 13) & 15u);
                np_rem = (int)(rng >> 17);
                kp = ks + ns_rem;
                ax = p.x - gx;
                ay = p.y - gy;
            }
            // primary records
            ikq_sts_u16<0>(wp, (unsigned)kp);
            ikq_sts_u16<2>(wp, (unsigned)(kp + 1));
            ikq_sts_u16<4>(wp, (unsigned)(kp + 2));
            ikq_sts_u16<6>(wp, (unsigned)(kp + 3));
            wp += 2u * (unsigned)min(max(np_rem, 0), 4);
            kp += 4;
            np_rem -= 4;
            // secondary records
            {
                const int nsl = min(max(ns_rem, 0), 2);
                const unsigned act = (1u << nsl) - 1u;
                const unsigned ba = b_addr + 16u * (unsigned)(nsl > 0 ? ks : 0);
                const double2 b0 = ikq_lds_f64x2<0>(ba), b1 = ikq_lds_f64x2<16>(ba);
                const double x0 = ax + b0.x, y0 = ay + b0.y, x1 = ax + b1.x, y1 = ay + b1.y;
                const double d0 = fma(x0, x0, y0 * y0), d1 = fma(x1, x1, y1 * y1);
                // kept when the primary end lies outside the disc, or inside the guard band of the test (phase 3 recomputes
                // the same distance as t(1), notices the band there and has the exact test decide)
                const unsigned keep = ((d0 > r2lo ? 1u : 0u) | (d1 > r2lo ? 2u : 0u)) & act;
                ikq_sts_u16<0>(wp, (unsigned)ks | IKQ_SEC);
                wp += (keep & 1u) << 1;
                ikq_sts_u16<0>(wp, (unsigned)(ks + 1) | IKQ_SEC);
                wp += keep & 2u;
                ks += 2;
                ns_rem -= 2;
            }
            const bool idle = ns_rem <= 0 && np_rem <= 0 && hp >= buf_top;
            state = (wp + 2u * IKQ_STEP_ROOM > hp) ? 2 : (idle ? 0 : 1);       // the next step's slots would reach my disc particles
            state = __reduce_max_sync(FULL_MASK, state);
        } while (state == 1);
        const int n = (int)((wp - buf_addr) >> 1);
        __syncwarp();

        // ---- phase 3: four lanes <-> grid point; eight points per pass, longest lists first so that the eight points
        //      of a pass have lists of similar length (the pass runs to the longest of them)
        int rank = 0;                                          // my list's rank by length (ties: by lane), 0 = longest
#pragma unroll
        for (int bit = 7; bit >= 0; --bit) {                  // list lengths are below 256 (ikq_pick_caps)
            // radix ranking from the top bit down: lanes with a larger (n, 31 - lane) key come first
            const unsigned ones = __ballot_sync(FULL_MASK, (n >> bit) & 1);
            const unsigned same = __match_any_sync(FULL_MASK, n >> (bit + 1));   // lanes that agree on the higher bits
            if (!((n >> bit) & 1)) rank += __popc(ones & same);
        }
        rank += __popc(__match_any_sync(FULL_MASK, n) & ((1u << lane) - 1u));   // equal lengths: lower lane first
        s_perm[(warp << 5) + rank] = (unsigned char)lane;
        __syncwarp();
#pragma unroll 1
        for (int pass = 0; pass < 4; ++pass) {
            // the point whose list has rank 8 pass + (lane >> 2) is my quad's
            const int pt = s_perm[(warp << 5) + pass * 8 + (lane >> 2)];
            const int n_p = __shfl_sync(FULL_MASK, n, pt);
            if (!__any_sync(FULL_MASK, n_p > 0)) break;        // the remaining lists are shorter still
            const double gxp = __shfl_sync(FULL_MASK, gx, pt), gyp = __shfl_sync(FULL_MASK, gy, pt);
            const unsigned lp_addr = buf_base + (unsigned)(((warp << 5) + pt) * caps.buf_stride) * 2u;
            double axx = 0.0, axy = 0.0, ayy = 0.0;
            unsigned unsure_any = 0u;
            if (sub < n_p) {
                // software pipeline: the next item's entry -> owner -> (position, b, (c2, f)) loads are issued before the
                // current item's quadrature, so the three dependent shared-memory round trips hide behind ~45 FP64 operations
                const unsigned a_last = lp_addr + 2u * (unsigned)(((n_p - 1 - sub) & ~3) + sub);   // my last entry
                unsigned a_it = lp_addr + 2u * (unsigned)sub;
                unsigned it0, it1;
                double2 p0, b0, cf0, p1, b1, cf1;
                auto fetch = [&](unsigned& it, double2& p, double2& b, double2& cf) {
                    it = ikq_lds_u16(a_it);
                    const unsigned kk = it & (unsigned)IKQ_MAX_NREC;
                    p = ikq_lds_f64x2<0>(pos_addr + 16u * ikq_lds_u16(own_addr + 2u * kk));
                    b = ikq_lds_f64x2<0>(b_addr + 16u * kk);
                    cf = ikq_lds_f64x2<0>(cf_addr + 16u * kk);
                };
                auto item = [&](unsigned it, const double2& p, const double2& b, const double2& cf) {
                    const double pax = p.x - gxp, pay = p.y - gyp;
                    const double c0 = fma(pax, pax, pay * pay) * iR2;
                    const double c1 = fma(pax, b.x, pay * b.y) * iR2x2;
                    // weight: a bond whose two ends lie in the disc is visited twice by the reference, 1/2 each time; the
                    // far end is inside when t(1) = |a + b|^2 / R^2 <= 1 (inside the guard band the pass is redone with
                    // the exact test)
                    const double t1 = (c0 + cf.x) + c1;
                    const bool out = t1 > t1hi;
                    const bool half = (it & IKQ_SEC) != 0u || out;
                    // inside the band, primary or secondary item alike (phase 2 keeps a secondary item only when its
                    // primary end is beyond R^2 - band / 2, i.e. beyond t1lo here, so none slips below the band)
                    unsure_any |= (!out && t1 > t1lo) ? 1u : 0u;
                    const double wf = ik_quadrature<KERNEL, Q>(A, c0, c1, cf.x) * cf.y;
                    const double wh = __hiloint2double(half ? 0x3FE00000 : 0x3FF00000, 0);
                    const double tx = (wf * wh) * b.x, ty = (wf * wh) * b.y;
                    axx = fma(tx, b.x, axx);
                    axy = fma(tx, b.y, axy);
                    ayy = fma(ty, b.y, ayy);
                };
                fetch(it0, p0, b0, cf0);
                for (;;) {                                         // two register sets, alternating: no copies
                    bool last = a_it >= a_last;
                    a_it = min(a_it + 8u, a_last);
                    fetch(it1, p1, b1, cf1);
                    item(it0, p0, b0, cf0);
                    if (last) break;
                    last = a_it >= a_last;
                    a_it = min(a_it + 8u, a_last);
                    fetch(it0, p0, b0, cf0);
                    item(it1, p1, b1, cf1);
                    if (last) break;
                }
            }
            if (__any_sync(FULL_MASK, unsure_any != 0u))                // rounding could matter for some far end: exact weights
                ikq_pass_exact<KERNEL, Q>(A, S, caps, lp_addr, n_p, sub, gxp, gyp, nruns, axx, axy, ayy);
            // quad reduction (same bits in all four lanes), then the sum goes to the point's own thread
            axx += __shfl_xor_sync(FULL_MASK, axx, 1); axy += __shfl_xor_sync(FULL_MASK, axy, 1); ayy += __shfl_xor_sync(FULL_MASK, ayy, 1);
            axx += __shfl_xor_sync(FULL_MASK, axx, 2); axy += __shfl_xor_sync(FULL_MASK, axy, 2); ayy += __shfl_xor_sync(FULL_MASK, ayy, 2);
            const int src = (rank & 7) << 2;                   // as owner of my point: its quad in pass rank >> 3
            const double vxx = __shfl_sync(FULL_MASK, axx, src), vxy = __shfl_sync(FULL_MASK, axy, src), vyy = __shfl_sync(FULL_MASK, ayy, src);
            if ((rank >> 3) == pass) { txx += vxx; txy += vxy; tyy += vyy; }
        }
        __syncwarp();
        if (state == 0) break;
    }

    if (store) {
        if (slow) ik_slow_path<KERNEL, Q>(A, gx, gy, txx, txy, tyy);
        *reinterpret_cast<double4*>(out_ptr) = make_double4(out_scale * txx, out_scale * txy, out_scale * txy, out_scale * tyy);
    }
}

static bool ik_quad_applicable(const PgtIKArgs& a)
{
    // rect stays with the two-phase kernel, which tests its nodes in the reference's exact operation order
    return a.grid.mode == 0 && a.obs_mask == PGT_OBS_VIRIAL && a.out_virial != nullptr && a.fr.n > 0 && a.fr.sxy != nullptr &&
           a.fr.hb_pm != nullptr && a.grid.span_x[0] >= 0.0 && a.cg.kernel != PGT_CG_RECT;
}

// Tile shape and shared-memory budget of one CTA from frame-wide statistics only (never from the shard), so that every
// shard of a grid runs the same code on the same tiles.  Returns false when no tile shape fits one SM.
static bool ikq_pick_caps(const PgtIKArgs& a, IKQCaps* caps, size_t* dyn_bytes)
{
    static int smem_sm = 0, smem_cta = 0;
    if (smem_sm == 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&smem_sm, cudaDevAttrMaxSharedMemoryPerMultiprocessor, dev);
        cudaDeviceGetAttribute(&smem_cta, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    }
    static int want_wx = -1, want_wy = 0, margin_pct = 10, want_list = 0, want_ctas = 0, slack_pct = 5;
    if (want_wx < 0) {
        const char* e = getenv("PGT_IKQ_TILE");              // "wx,wy"
        want_wx = 0;
        if (e) sscanf(e, "%d,%d", &want_wx, &want_wy);
        if ((e = getenv("PGT_IKQ_MARGIN"))) margin_pct = atoi(e);
        if ((e = getenv("PGT_IKQ_LIST"))) want_list = atoi(e);
        if ((e = getenv("PGT_IKQ_CTAS"))) want_ctas = atoi(e);
        if ((e = getenv("PGT_IKQ_SLACK"))) slack_pct = atoi(e);
    }
    const double R = a.cg.rcut, cwx = 1.0 / a.cells.inv_cwx, cwy = 1.0 / a.cells.inv_cwy;
    // Every capacity below comes from quantities that do not depend on the grid shard -- the frame's particle count as
    // uploaded, the box, the cell geometry, the grid's tile extents -- never from the particles or bonds a slab-culled
    // frame happens to hold: a point's summation order (its rounds, its path) must be the same on 1 GPU and on 8.
    const double rho = (double)std::max<long long>(a.n_frame, 1) / (a.box.L[0] * a.box.L[1]);
    // half-bonds per particle, estimated: a share of the neighbours within max_rcut (0.29 on polydisperse liquids: the
    // pair cut-off scales with the pair's diameters, the search radius with the largest); it sizes the lists only,
    // the record store takes whatever shared memory is left
    const double hb = std::max(1.0, 0.29 * rho * M_PI * a.max_rcut * a.max_rcut);
    const double m = 1.0 + 0.01 * margin_pct;
    static const int shapes[][2] = { { 4, 2 }, { 2, 4 }, { 2, 2 }, { 1, 2 }, { 1, 1 } };   // equal threads: the first listed wins (fewer, longer cell rows)
    const int reg_threads = 65536 / 128;                       // __launch_bounds__(256, 2): at most 128 registers per thread
    long long best_score = -1;
    double best_mt = 0.0;
    // a thread's buffer holds the particles of its disc (mean rho pi R^2; 1.6 x that + 8, a fuller disc takes the slow
    // path) and, growing towards them, the bonds to integrate (about 2/3 of the disc's half-bonds): room for a whole
    // list saves a second walk / integrate round
    const double nd = rho * M_PI * R * R;
    const int buf_min = (((int)(1.6 * nd) + 8 + IKQ_STEP_ROOM + 2) + 1) & ~1;
    if (buf_min > 250) return false;                           // phase 3 ranks the lists by an 8-bit length
    const int buf_want = std::max(buf_min, std::min(64, (((int)(0.66 * 1.5 * nd * hb) + IKQ_STEP_ROOM + 4) + 1) & ~1));
    for (const auto& sh : shapes) {
        const int wx = sh[0], wy = sh[1];
        if (want_wx > 0 && (wx != want_wx || wy != want_wy)) continue;
        const int T = 32 * wx * wy;
        const double sx = a.grid.span_x[wx - 1], sy = a.grid.span_y[wy - 1];
        if (sx < 0.0 || sy < 0.0) continue;
        const int rows_cap = (int)((sy + 2.0 * R) / cwy) + 3, cols = (int)((sx + 2.0 * R) / cwx) + 3;
        if (rows_cap > IKQ_MAXROWS || cols > 254) continue;
        const int lcs_stride = cols + 1;
        const double ep = rho * (sx + 2.0 * R + 2.0 * cwx) * (sy + 2.0 * R + 2.0 * cwy);
        const long long nloc = ((long long)(ep * m + 32.0) + 7) & ~7LL;
        const long long nrec_want = ((long long)(ep * hb * m + 64.0) + 7) & ~7LL;    // records of a rectangle, with margin
        const long long nrec_need = ((long long)(ep * hb * 1.05 + 32.0) + 7) & ~7LL;  // fewer than this: tiles would overflow
        if (nloc > 0x7FFF) continue;
        const long long lc_top = want_list > 0 ? std::min(254, std::max(want_list, buf_min)) & ~1 : buf_want;
        const long long stat = (long long)sizeof(IKQStatic) + 16 + IKQ_MAX_THREADS + 1024;   // static part + the per-CTA reservation
        const int max_ctas = std::max(1, std::min(reg_threads / T, 16));
        bool placed = false;
        for (int ctas = (want_ctas > 0 ? std::min(want_ctas, max_ctas) : max_ctas); ctas >= 1 && !placed; --ctas) {
            // the longest per-thread buffer (down to 8 entries below the wish) that leaves room for the records
            for (long long lc = lc_top; lc >= std::max<long long>(buf_min, lc_top - 8) && !placed; lc -= 2) {
                const long long stride = ((lc / 2) | 1) * 2;   // lc is even: the next 2 x odd, lc or lc + 2
                // position 16 B + 8 B (single precision), range 4 B per particle; local cell starts; the threads' buffers
                const long long base = nloc * 28 + (long long)rows_cap * lcs_stride * 2 + stride * 2 * T + 64;
                // records 32 B + owner 2 B: the shared memory that is left, up to a slack over the wish (the rest stays L1)
                const long long avail = std::min<long long>(smem_sm / ctas, smem_cta) - stat - base - 16;
                const long long nrec = std::min<long long>(std::min<long long>((avail / 34) & ~7LL, IKQ_MAX_NREC & ~7),
                                                          (nrec_want * (100 + slack_pct) / 100 + 7) & ~7LL);
                if (nrec < nrec_need) continue;
                placed = true;
                const long long score = (long long)ctas * T;           // threads per SM; ties: the shape listed first
                if (score > best_score) {
                    best_score = score;
                    best_mt = std::max(sx, sy) + R + 2.0 * std::max(cwx, cwy);
                    caps->wx = wx; caps->wy = wy; caps->wy_log2 = wy == 4 ? 2 : (wy == 2 ? 1 : 0);
                    caps->rows_cap = rows_cap;
                    caps->lcs_stride = lcs_stride;
                    caps->nloc_cap = (int)nloc;
                    caps->nrec_cap = (int)nrec;
                    caps->buf_cap = (int)lc;
                    caps->buf_stride = (int)stride;
                    caps->rank_bits = lc < 64 ? 6 : 8;
                    *dyn_bytes = (size_t)(base + nrec * 34 + 15) & ~(size_t)15;
                }
            }
        }
    }
    if (best_score < 0) return false;
    // guard band of the arithmetic far-end test: a + b carries a few roundings at coordinate magnitude M
    const double M = 0.5 * std::max(a.box.L[0], a.box.L[1]) + R;
    const double band = std::max(1e-9 * a.cg.rcutsq, 1e-11 * R * M);
    caps->r2lo = a.cg.rcutsq - band;
    caps->r2hi = a.cg.rcutsq + band;
    caps->r2keep = a.cg.rcutsq - 0.5 * band;
    caps->t1lo = 1.0 - band * a.inv_rcutsq;
    caps->t1hi = 1.0 + band * a.inv_rcutsq;
    // guard band of phase 1's single-precision test: coordinates relative to the tile's corner are at most Mt in
    // magnitude, each rounded to 2^-24 relative; d^2 = dx^2 + dy^2 near R^2 then carries at most
    // 2^-24 (2.83 R (2 Mt + R + 1) + 3 R^2); twice that on either side
    const double e32 = std::ldexp(1.0, -23) * (2.83 * R * (2.0 * best_mt + R + 1.0) + 3.0 * R * R);
    caps->r2lo32 = std::nextafter((float)(a.cg.rcutsq - e32), -1.0f);
    caps->r2hi32 = std::nextafter((float)(a.cg.rcutsq + e32), 1e30f);
    return true;
}

template <int KERNEL, int Q>
static cudaError_t launch_ik_quad_t(const PgtIKArgs& a, const IKQCaps& caps, size_t dyn, cudaStream_t st)
{
    if (a.grid.g1 <= a.grid.g0) return cudaSuccess;
    static size_t attr_dyn_dev[64] = { 0 };                 // one per template instance and device (the attribute is per device)
    int dev = 0;
    cudaGetDevice(&dev);
    size_t& attr_dyn = attr_dyn_dev[dev & 63];
    if (dyn > attr_dyn) {
        cudaError_t e = cudaFuncSetAttribute(k_ik_quad<KERNEL, Q>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn);
        if (e != cudaSuccess) return e;
        cudaFuncSetAttribute(k_ik_quad<KERNEL, Q>, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared);
        attr_dyn = dyn;
    }
    const int TI = IKQ_PX * caps.wx, TJ = IKQ_PY * caps.wy;
    const int tiles_j = (a.grid.ny + TJ - 1) / TJ;
    const int i_lo = (int)(a.grid.g0 / a.grid.ny), i_hi = (int)((a.grid.g1 - 1) / a.grid.ny);
    const int tile_i_begin = i_lo / TI;
    const dim3 blocks((unsigned)tiles_j, (unsigned)(i_hi / TI - tile_i_begin + 1));
    if (blocks.y > 65535u) return cudaErrorInvalidConfiguration;
    k_ik_quad<KERNEL, Q><<<blocks, 32 * caps.wx * caps.wy, dyn, st>>>(a, tile_i_begin, caps);
    return cudaGetLastError();
}

template <int KERNEL>
static cudaError_t launch_ik_quad_q(const PgtIKArgs& a, const IKQCaps& caps, size_t dyn, cudaStream_t st)
{
    if (a.nq == 5) return launch_ik_quad_t<KERNEL, 5>(a, caps, dyn, st);
    return launch_ik_quad_t<KERNEL, 0>(a, caps, dyn, st);
}

static cudaError_t launch_ik_quad(const PgtIKArgs& a, const IKQCaps& caps, size_t dyn, cudaStream_t st)
{
    switch (a.cg.kernel) {
    case PGT_CG_OCTIC: return launch_ik_quad_q<PGT_CG_OCTIC>(a, caps, dyn, st);
    default: return launch_ik_quad_q<PGT_CG_MOLLIFIER>(a, caps, dyn, st);
    }
}
