// pgt_ik_walk.inl -- stage 4, Cartesian grids: the "walk" gather (included by pgt_kernels.cu after pgt_ik_tile.inl).
//
// A CTA of 128 threads owns a 16 x 8 tile of grid points; a thread owns ONE grid point from start to finish, so there is
// no cross-lane bookkeeping (no prefix sums, owner tables or shuffle reductions) and no CTA barrier after staging.
//
//  stage     The cells within R of the tile form a rectangle of cell rows; the frame is sorted by cell (row-major), so
//            every row of the rectangle is ONE contiguous run of particles and ONE contiguous run of half-bond records
//            (two when the tile straddles the periodic seam).  A warp takes a run: positions, each particle's record
//            range, each record's owner and the tile-local index of its far end j (found by counting the row starts
//            at or below j; bit 15 marks j < i) go to shared memory with coalesced loads.  The records themselves
//            (b, (c2, f)) are read through L1 by default; STAGE = true copies them to shared memory as well.
//  phase 1   Each thread tests the particles of ITS cell window exactly as Aboria does (dx = r_p - (x + kL),
//            dx^2 + dy^2 <= R^2); the hits of a window row are collected in a register mask and OR-ed into a private
//            bitmap over tile-local particle indices.
//  phase 2   "walk": one step = the next particle i of my disc (if the current one is exhausted), then four of its
//            half-bonds i -> j.  A bond whose far end j is also in the disc (one bitmap probe) is visited twice by the
//            reference, from i and from j, each time with weight 1/2 and the same line integral: it is kept ONCE (from
//            the end with the smaller local index) with weight 1; a bond with only i inside keeps weight 1/2.  That is
//            the reference's sum (IrvingKirkwood.h:80-107) with U(x) quadratures instead of ~1.5 U(x).  Survivors go
//            to a private list in shared memory ([slot][thread], 2 bytes: record + weight bit).
//  phase 3   Each thread integrates its own list: a = r_i - x, t(s) = c0 + s (c1 + s c2), Gauss-Legendre sum of the
//            shape, 3 FMAs into register accumulators.  Phases 2 and 3 alternate until the walk is finished (the list
//            holds `list_cap` entries per thread), on warp-uniform conditions only, so the warps of a CTA drift apart
//            and overlap integer and FP64 work.
//
// Everything discrete (disc membership, bond set, weights) is bit-identical to the reference arithmetic; a grid
// point's items are visited in an order that depends on nothing but the frame and the point, so the field does not
// depend on the sharding.  A CTA whose tile does not meet the limits runs `ikt_slow_path` (plain half-bond
// formulation from global memory): same results, no limits.
#define IKW_TI 16
#define IKW_TJ 8
#define IKW_THREADS (IKW_TI * IKW_TJ)
#define IKW_WARPS (IKW_THREADS / 32)
#define IKW_MAXROWS 32
#define IKW_MAXRUNS (2 * IKW_MAXROWS)
#define IKW_MAXCOLS 62
#define IKW_BOTH 0x4000u
#define IKW_MAX_NREC 0x3FFF              // a bond entry addresses 14 bits
#define IKW_MAX_NLOC 0x7FC0              // a far-end entry addresses 15 bits

struct IKWCaps {
    int rows_cap;                        // cell rows of the rectangle
    int nloc_cap;                        // particles of the rectangle (multiple of 32)
    int nrec_cap;                        // half-bond records of the rectangle
    int list_cap;                        // list entries per thread
};

struct IKWStatic {
    int ps[IKW_MAXRUNS];                 // first sorted particle of the run
    int np[IKW_MAXRUNS];                 // particles in the run
    int h0[IKW_MAXRUNS];                 // first half-bond record of the run (global)
    int lp_base[IKW_MAXRUNS + 1];        // tile-local index of the run's first particle
    int lr_base[IKW_MAXRUNS + 1];        // tile-local index of the run's first record
    double2 shift[IKW_MAXRUNS];          // (kx Lx, ky Ly) of the run's periodic image
};

// Gauss-Legendre sum of the shape along the bond.  5-point octic: g = 1 - t^2 is positive exactly when the node lies
// inside the disc (t < 1), so the reference's `if (d^2 < R^2)` is a clamp of g at zero -- done on the sign word with two
// integer instructions instead of a compare and two selects (a cleared high word leaves at most a denormal, whose
// square is zero).
template <int KERNEL, int Q>
__device__ __forceinline__ double ikw_quadrature(const PgtIKArgs& A, double c0, double c1, double c2)
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
    return bond_quadrature<KERNEL, Q>(c0, c1, c2, A.nq);
}

template <int KERNEL, int Q, bool STAGE>
__global__ void __launch_bounds__(IKW_THREADS) k_ik_walk(const PgtIKArgs A, const int tile_i_begin, const IKWCaps caps)
{
    __shared__ IKWStatic S;
    extern __shared__ __align__(16) unsigned char ikw_dyn[];
    // STAGE: the rectangle's records live in shared memory; else they are read through L1 (more CTAs per SM)
    double2* const s_recb = reinterpret_cast<double2*>(ikw_dyn);
    double2* const s_reccf = s_recb + (STAGE ? caps.nrec_cap : 0);
    double2* const s_pos = s_reccf + (STAGE ? caps.nrec_cap : 0);
    unsigned* const s_bits = reinterpret_cast<unsigned*>(s_pos + caps.nloc_cap);        // [word][thread], word NW stays zero
    const int NW = caps.nloc_cap >> 5;
    unsigned* const s_rng = s_bits + (NW + 1) * IKW_THREADS;                            // k0 | kend << 16 per particle
    int* const s_gd = reinterpret_cast<int*>(s_rng + caps.nloc_cap);                    // !STAGE: global minus local record index
    unsigned short* const s_list = reinterpret_cast<unsigned short*>(s_gd + (STAGE ? 0 : caps.nloc_cap));   // [slot][thread]
    unsigned short* const s_jl = s_list + caps.list_cap * IKW_THREADS;                  // local index of j | (j < i) << 15
    unsigned short* const s_own = s_jl + caps.nrec_cap;                                 // local index of i
    unsigned short* const s_lcs = s_own + caps.nrec_cap;                                // [row][IKW_MAXCOLS + 2]: local index of the first particle of a cell
    const unsigned NOJ = (unsigned)caps.nloc_cap;                                       // its bitmap word is the zero word NW

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const PgtFrameView& fr = A.fr;
    const double R = A.cg.rcut, R2 = A.cg.rcutsq, iR2 = A.inv_rcutsq;
    const int ncx = A.cells.ncx, ncy = A.cells.ncy;

    // ---- tile and my grid point: a warp covers a compact 8 x 4 patch
    const int tiles_j = (A.grid.ny + IKW_TJ - 1) / IKW_TJ;
    const int i0 = (tile_i_begin + blockIdx.x / tiles_j) * IKW_TI, j0 = (blockIdx.x % tiles_j) * IKW_TJ;
    const int i1 = min(i0 + IKW_TI, A.grid.nx) - 1, j1 = min(j0 + IKW_TJ, A.grid.ny) - 1;
    const int gi = i0 + 8 * (warp >> 1) + (lane >> 2), gj = j0 + 4 * (warp & 1) + (lane & 3);
    bool valid = gi <= i1 && gj <= j1;
    long long g = 0;
    double gx = 0.0, gy = 0.0;
    if (valid) {
        g = (long long)gi * A.grid.ny + gj;
        valid = g >= A.grid.g0 && g < A.grid.g1;
        gx = __ldg(A.grid.xs + gi);
        gy = __ldg(A.grid.ys + gj);
    }

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
    // static limits; the rectangle plus a bond's reach must not wrap onto itself (one image of every particle)
    bool fast = ncols <= IKW_MAXCOLS && nrows <= caps.rows_cap && ncols + 2 * A.cells.reach + 1 <= ncx &&
                nrows + 2 * A.cells.reach + 1 <= ncy && ux0 >= -ncx && ux1 < 2 * ncx && uy0 >= -ncy && uy1 < 2 * ncy;

    Seg seg[3];
    int nseg = 0;
    if (fast) nseg = split_segments(ux0, ux1, ncx, seg);
    if (nseg > 2) fast = false;
    const int nruns = 2 * nrows;
    const int csplit = (nseg == 2) ? (seg[0].c1 - seg[0].c0 + 1) : ncols;   // column at which the second segment starts
    const bool seam = fast && (nseg == 2 || seg[0].k != 0 || uy0 < 0 || uy1 >= ncy);   // some run carries an image shift

    int nloc = 0, nrec_total = 0;
    if (fast) {
        // ---- runs: (row, segment) -> contiguous particles [ps, ps+np) and records
        if (tid < nruns) {
            const int r = tid >> 1, s = tid & 1;
            int uy = uy0 + r, ky = 0;
            if (uy < 0) { ky = +1; uy += ncy; } else if (uy >= ncy) { ky = -1; uy -= ncy; }
            int ps = 0, pe = 0, h0 = 0, h1 = 0, kx = 0;
            if (s < nseg) {
                ps = __ldg(fr.cell_start + uy * ncx + seg[s].c0);
                pe = __ldg(fr.cell_start + uy * ncx + seg[s].c1 + 1);
                kx = seg[s].k;
                if (pe > ps) { h0 = __ldg(fr.hb_off + ps); h1 = __ldg(fr.hb_off + pe); }
            }
            S.ps[tid] = ps;
            S.np[tid] = pe - ps;
            S.h0[tid] = h0;
            S.lr_base[tid] = h1 - h0;                      // counts for now, scanned below
            S.shift[tid] = make_double2(kx * A.box.L[0], ky * A.box.L[1]);
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
        if (nloc > caps.nloc_cap || nrec_total > caps.nrec_cap) fast = false;     // uniform: every thread reads the same totals
    }

    double txx = 0.0, txy = 0.0, tyy = 0.0;
    const double out_scale = A.cg.norm;                    // weights 1 / 0.5 already carry VirialStress::compute_cg's 1/2
    double* const out_ptr = A.out_virial + 4 * (g - A.grid.g0);

    if (!fast) {
        if (valid) {
            ikt_slow_path<KERNEL, Q>(A, gx, gy, txx, txy, tyy);
            *reinterpret_cast<double4*>(out_ptr) = make_double4(out_scale * txx, out_scale * txy, out_scale * txy, out_scale * tyy);
        }
        return;
    }

    // ---- local cell starts
    for (int e = tid; e < nrows * (ncols + 1); e += IKW_THREADS) {
        const int r = e / (ncols + 1), c = e - r * (ncols + 1);
        int v;
        if (c == ncols) v = S.lp_base[2 * r + 2];
        else {
            int uy = uy0 + r;
            if (uy < 0) uy += ncy; else if (uy >= ncy) uy -= ncy;
            const int s = c >= csplit;
            const int cell = uy * ncx + (s ? seg[1].c0 + (c - csplit) : seg[0].c0 + c);
            v = S.lp_base[2 * r + s] + (__ldg(fr.cell_start + cell) - S.ps[2 * r + s]);
        }
        s_lcs[r * (IKW_MAXCOLS + 2) + c] = (unsigned short)v;
    }
    // ---- a warp takes a run: particles, then the run's records
    for (int run = warp; run < nruns; run += IKW_WARPS) {
        const int ps = S.ps[run], np = S.np[run], lpb = S.lp_base[run], h0 = S.h0[run], lrb = S.lr_base[run];
        for (int o = lane; o < np; o += 32) {
            const int p = ps + o, l = lpb + o;
            s_pos[l] = make_double2(__ldg(fr.sx + p), __ldg(fr.sy + p));
            const int k0 = lrb + (__ldg(fr.hb_off + p) - h0), k1 = lrb + (__ldg(fr.hb_off + p + 1) - h0);
            s_rng[l] = (unsigned)k0 | ((unsigned)k1 << 16);
            if (!STAGE) s_gd[l] = h0 - lrb;
            for (int k = k0; k < k1; ++k) s_own[k] = (unsigned short)l;
        }
        __syncwarp();
        const int nr = S.lr_base[run + 1] - lrb;
        // j lies within `reach` cell rows of i: probe those rows' runs for the one that holds it
        const int r = run >> 1;
        const int q0 = 2 * max(r - A.cells.reach, 0), q1 = 2 * min(r + A.cells.reach, nrows - 1) + 1;
        for (int o = lane; o < nr; o += 32) {
            const int e = h0 + o;
            if (STAGE) {
                s_recb[lrb + o] = __ldg(fr.hb_b + e);
                s_reccf[lrb + o] = __ldg(fr.hb_cf + e);
            }
            const int j = __ldg(fr.hb_j + e);
            unsigned jl = NOJ;
            if (!seam) {
                // one run per row, ascending with the row: count the row starts at or below j, no divergent probing
                int rr = q0 >> 1;
                for (int q = (q0 >> 1) + 1; q <= (q1 >> 1); ++q) rr += (j >= S.ps[2 * q]) ? 1 : 0;
                const unsigned off = (unsigned)(j - S.ps[2 * rr]);
                if (off < (unsigned)S.np[2 * rr]) jl = (unsigned)S.lp_base[2 * rr] + off;
            } else {
                for (int q = q0; q <= q1; ++q) {
                    const unsigned off = (unsigned)(j - S.ps[q]);
                    if (off < (unsigned)S.np[q]) { jl = (unsigned)S.lp_base[q] + off; break; }
                }
            }
            // the bond is integrated from its end with the smaller local index when both ends are in a disc
            if (jl != NOJ && jl < (unsigned)s_own[lrb + o]) jl |= 0x8000u;
            s_jl[lrb + o] = (unsigned short)jl;
        }
    }
    unsigned* const mybits = s_bits + tid;
    for (int w = 0; w <= NW; ++w) mybits[w * IKW_THREADS] = 0u;
    __syncthreads();

    // ================= phase 1: my disc as a bitmap over tile-local particle indices =================
    int w_first = NW, w_last = -1;
    if (valid) {
        const double padx = 1e-9 * (R + fabs(gx) + 1.0), pady = 1e-9 * (R + fabs(gy) + 1.0);
        const int cx0 = max(__double2int_rd((gx - R - padx - A.box.lo[0]) * A.cells.inv_cwx), ux0) - ux0;
        const int cx1 = min(__double2int_rd((gx + R + padx - A.box.lo[0]) * A.cells.inv_cwx), ux1) - ux0;
        const int ry0 = max(__double2int_rd((gy - R - pady - A.box.lo[1]) * A.cells.inv_cwy), uy0) - uy0;
        const int ry1 = min(__double2int_rd((gy + R + pady - A.box.lo[1]) * A.cells.inv_cwy), uy1) - uy0;
        // a window row's candidates are consecutive local indices: collect the hits of (at most) 32 of them in a
        // register mask, then OR the mask into the one or two bitmap words it straddles
        auto scan_rows = [&](auto SEAM) {
            for (int r = ry0; r <= ry1; ++r) {
                const int la = s_lcs[r * (IKW_MAXCOLS + 2) + cx0], lb = s_lcs[r * (IKW_MAXCOLS + 2) + cx1 + 1];
                double cpy = gy, cpxA = gx, cpxB = gx;
                int lsplit = lb;
                if (SEAM.value) {                              // x + k (bmax - bmin), Search.h:186-189
                    const double2 shA = S.shift[2 * r], shB = S.shift[2 * r + 1];
                    cpy = gy + shA.y; cpxA = gx + shA.x; cpxB = gx + shB.x;
                    lsplit = S.lp_base[2 * r + 1];
                }
                for (int lc = la; lc < lb; lc += 32) {
                    const int le = min(lb, lc + 32);
                    unsigned mask = 0u;
                    for (int l = lc; l < le; ++l) {
                        const double2 p = s_pos[l];
                        const double ax = p.x - ((SEAM.value && l >= lsplit) ? cpxB : cpxA);
                        const double ay = p.y - cpy;
                        const double d2 = ax * ax + ay * ay;
                        mask |= (d2 > R2 ? 0u : 1u) << (l - lc);
                    }
                    if (mask) {
                        const int w0 = lc >> 5, sh = lc & 31;
                        mybits[w0 * IKW_THREADS] |= mask << sh;
                        w_first = min(w_first, w0);
                        int wl = w0;
                        if (sh != 0 && (mask >> (32 - sh)) != 0u) { mybits[(w0 + 1) * IKW_THREADS] |= mask >> (32 - sh); wl = w0 + 1; }
                        w_last = max(w_last, wl);
                    }
                }
            }
        };
        if (seam) scan_rows(std::true_type{}); else scan_rows(std::false_type{});
    }
    if (seam) {
        // positions of the tile's periodic image: r - kL (used for the quadrature only, never for a membership test)
        __syncthreads();
        for (int run = warp; run < nruns; run += IKW_WARPS) {
            const int np = S.np[run], lpb = S.lp_base[run];
            const double2 sh = S.shift[run];
            for (int o = lane; o < np; o += 32) {
                const double2 p = s_pos[lpb + o];
                s_pos[lpb + o] = make_double2(p.x - sh.x, p.y - sh.y);
            }
        }
        __syncthreads();
    }

    // ================= phases 2 and 3, alternating; warp-uniform control only =================
    unsigned short* const mylist = s_list + tid;
    const int C = caps.list_cap;
    int w = w_first;
    unsigned bits = (valid && w <= w_last) ? mybits[w * IKW_THREADS] : 0u;
    if (!valid) w_last = -1;
    int k = 0, kend = 0;
    const double iR2x2 = 2.0 * iR2;

    for (;;) {
        // ---- phase 2: walk.  One step = (if my particle is exhausted) the next particle of my disc, then four of its
        //      half-bonds; a vote per step: stop when some list is nearly full or every lane has finished
        int n = 0, state;
        do {
            if (k >= kend) {
                if (bits == 0u && w < w_last) { ++w; bits = mybits[w * IKW_THREADS]; }
                if (bits) {
                    const int l = (w << 5) + __ffs(bits) - 1;
                    bits &= bits - 1u;
                    const unsigned rng = s_rng[l];
                    k = (int)(rng & 0xFFFFu);
                    kend = (int)(rng >> 16);
                }
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                if (k + u < kend) {
                    const unsigned e = s_jl[k + u], jl = e & 0x7FFFu;
                    const unsigned both = (mybits[(jl >> 5) * IKW_THREADS] >> (jl & 31u)) & 1u;
                    if (!(both & (e >> 15))) {                     // else: kept from j's side
                        mylist[n * IKW_THREADS] = (unsigned short)((unsigned)(k + u) | (both << 14));
                        ++n;
                    }
                }
            }
            k += 4;
            const bool idle = k >= kend && bits == 0u && w >= w_last;
            state = (n > C - 4) ? 2 : (idle ? 0 : 1);
            state = __reduce_max_sync(FULL_MASK, state);
        } while (state == 1);

        // ---- phase 3: integrate my list (the next entry and its owner are fetched one item ahead)
        unsigned it = 0u, ow = 0u;
        if (n > 0) { it = mylist[0]; ow = s_own[it & 0x3FFFu]; }
        for (int s = 0; s < n; ++s) {
            const unsigned kk = it & 0x3FFFu, cur_ow = ow;
            const double wgt = (it & IKW_BOTH) ? 1.0 : 0.5;
            if (s + 1 < n) { it = mylist[(s + 1) * IKW_THREADS]; ow = s_own[it & 0x3FFFu]; }
            const double2 p = s_pos[cur_ow];
            double2 b, cf;
            if (STAGE) { b = s_recb[kk]; cf = s_reccf[kk]; }
            else { const int e = (int)kk + s_gd[cur_ow]; b = __ldg(fr.hb_b + e); cf = __ldg(fr.hb_cf + e); }
            const double ax = p.x - gx, ay = p.y - gy;
            const double c0 = fma(ax, ax, ay * ay) * iR2;
            const double c1 = fma(ax, b.x, ay * b.y) * iR2x2;
            const double wf = ikw_quadrature<KERNEL, Q>(A, c0, c1, cf.x) * (cf.y * wgt);
            const double tx = wf * b.x, ty = wf * b.y;
            txx = fma(tx, b.x, txx);
            txy = fma(tx, b.y, txy);
            tyy = fma(ty, b.y, tyy);
        }
        if (state == 0) break;
    }

    if (valid) *reinterpret_cast<double4*>(out_ptr) = make_double4(out_scale * txx, out_scale * txy, out_scale * txy, out_scale * tyy);
}

static bool ik_walk_applicable(const PgtIKArgs& a)
{
    return a.grid.mode == 0 && a.obs_mask == PGT_OBS_VIRIAL && a.out_virial != nullptr && a.fr.n > 0 && a.grid.tspan_x >= 0.0;
}

// Shared-memory budget of one CTA from frame-wide statistics only (never from the shard), so that every shard of a
// grid runs the same code on the same tiles.  Returns false when the rectangle cannot fit one SM.
static bool ikw_pick_caps(const PgtIKArgs& a, bool stage, IKWCaps* caps, size_t* dyn_bytes)
{
    static int smem_sm = 0, smem_cta = 0;
    if (smem_sm == 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&smem_sm, cudaDevAttrMaxSharedMemoryPerMultiprocessor, dev);
        cudaDeviceGetAttribute(&smem_cta, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    }
    static int want_ctas = -1, margin_pct = 15, want_list = 0;
    if (want_ctas < 0) {
        const char* e = getenv("PGT_IKW_CTAS");
        want_ctas = e ? atoi(e) : 0;
        if ((e = getenv("PGT_IKW_MARGIN"))) margin_pct = atoi(e);
        if ((e = getenv("PGT_IKW_LIST"))) want_list = atoi(e);
    }
    const double R = a.cg.rcut, cwx = 1.0 / a.cells.inv_cwx, cwy = 1.0 / a.cells.inv_cwy;
    const double area = (a.grid.tspan_x + 2.0 * R + 2.0 * cwx) * (a.grid.tspan_y + 2.0 * R + 2.0 * cwy);
    const double rho = (double)a.fr.n / (a.box.L[0] * a.box.L[1]);
    const double hb = a.fr.n_hb > 0 ? (double)a.fr.n_hb / (double)a.fr.n : 1.0;
    const double m = 1.0 + 0.01 * margin_pct;
    const double ep = rho * area;
    const long long nloc = ((long long)(ep * m + 32.0) + 31) & ~31LL;
    const long long nrec = ((long long)(ep * hb * m + 64.0) + 7) & ~7LL;
    if (nloc > IKW_MAX_NLOC || nrec > IKW_MAX_NREC) return false;
    // records 32 B (staged), far-end and owner tables 2 + 2 B per record; position 16 B, range 4 B (+ 4 B offset) per
    // particle; one bitmap word per 32 particles (+ the zero word) per thread
    const int rows_cap = std::min<int>(IKW_MAXROWS, (int)((a.grid.tspan_y + 2.0 * R) / cwy) + 3);
    const long long base = nrec * ((stage ? 32 : 0) + 4) + nloc * (stage ? 20 : 24) + (nloc / 32 + 1) * IKW_THREADS * 4 +
                           (long long)rows_cap * (IKW_MAXCOLS + 2) * 2 + 64;
    const long long stat = (long long)sizeof(IKWStatic) + 1024;          // static part + the per-CTA reservation
    // the most CTAs per SM that still leave every thread a list of >= 32 entries (>= 8 when a count was asked for)
    int list_cap = 0;
    for (int ctas = (want_ctas > 0 ? want_ctas : 8); ctas >= 1 && list_cap == 0; --ctas) {
        const long long avail = std::min<long long>(smem_sm / ctas, smem_cta) - stat - base;
        const long long lc = avail / (2 * IKW_THREADS);
        const long long need = (ctas == 1 || ctas == want_ctas) ? 8 : 32;
        if (lc >= need) list_cap = (int)std::min<long long>(lc, want_list > 0 ? want_list : 32);
    }
    if (list_cap == 0) return false;
    caps->rows_cap = rows_cap;
    caps->nloc_cap = (int)nloc;
    caps->nrec_cap = (int)nrec;
    caps->list_cap = list_cap;
    *dyn_bytes = (size_t)(base + (long long)list_cap * 2 * IKW_THREADS + 15) & ~(size_t)15;
    return true;
}

template <int KERNEL, int Q, bool STAGE>
static cudaError_t launch_ik_walk_t(const PgtIKArgs& a, const IKWCaps& caps, size_t dyn, cudaStream_t st)
{
    if (a.grid.g1 <= a.grid.g0) return cudaSuccess;
    static size_t attr_dyn_dev[64] = { 0 };                 // one per template instance and device (the attribute is per device)
    int dev = 0;
    cudaGetDevice(&dev);
    size_t& attr_dyn = attr_dyn_dev[dev & 63];
    if (dyn > attr_dyn) {
        cudaError_t e = cudaFuncSetAttribute(k_ik_walk<KERNEL, Q, STAGE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn);
        if (e != cudaSuccess) return e;
        const char* ce = getenv("PGT_IKW_CARVE");
        cudaFuncSetAttribute(k_ik_walk<KERNEL, Q, STAGE>, cudaFuncAttributePreferredSharedMemoryCarveout,
                             ce ? atoi(ce) : (int)cudaSharedmemCarveoutMaxShared);
        attr_dyn = dyn;
    }
    const int tiles_j = (a.grid.ny + IKW_TJ - 1) / IKW_TJ;
    const int i_lo = (int)(a.grid.g0 / a.grid.ny), i_hi = (int)((a.grid.g1 - 1) / a.grid.ny);
    const int tile_i_begin = i_lo / IKW_TI;
    const long long blocks = (long long)(i_hi / IKW_TI - tile_i_begin + 1) * tiles_j;
    k_ik_walk<KERNEL, Q, STAGE><<<(unsigned)blocks, IKW_THREADS, dyn, st>>>(a, tile_i_begin, caps);
    return cudaGetLastError();
}

template <int KERNEL>
static cudaError_t launch_ik_walk_q(const PgtIKArgs& a, const IKWCaps& caps, size_t dyn, bool stage, cudaStream_t st)
{
    if (a.nq == 5) return stage ? launch_ik_walk_t<KERNEL, 5, true>(a, caps, dyn, st) : launch_ik_walk_t<KERNEL, 5, false>(a, caps, dyn, st);
    return stage ? launch_ik_walk_t<KERNEL, 0, true>(a, caps, dyn, st) : launch_ik_walk_t<KERNEL, 0, false>(a, caps, dyn, st);
}

static cudaError_t launch_ik_walk(const PgtIKArgs& a, const IKWCaps& caps, size_t dyn, bool stage, cudaStream_t st)
{
    switch (a.cg.kernel) {
    case PGT_CG_OCTIC: return launch_ik_walk_q<PGT_CG_OCTIC>(a, caps, dyn, stage, st);
    case PGT_CG_RECT: return launch_ik_walk_q<PGT_CG_RECT>(a, caps, dyn, stage, st);
    default: return launch_ik_walk_q<PGT_CG_MOLLIFIER>(a, caps, dyn, stage, st);
    }
}
