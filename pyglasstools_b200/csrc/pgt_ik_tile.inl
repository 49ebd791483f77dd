// pgt_ik_tile.inl -- stage 4, Cartesian grids: the "tile" gather (included by pgt_kernels.cu).
//
// A CTA of 256 threads owns a 16 x 16 tile of grid points.  No global atomics, fixed summation order.
//
//  stage     The cells within R of the tile form a rectangle of cell rows; because the frame is sorted by cell
//            (row-major), every row of the rectangle is ONE contiguous run of particles in HBM (two when the tile
//            straddles the periodic seam) and ONE contiguous run of half-bond records.  Stored positions, record
//            offsets and -- translated once per CTA -- the tile-local index of every bond's j go to shared memory.
//  phase 1   thread <-> grid point.  Each thread tests the particles of ITS cell window exactly as Aboria does
//            (dx = r_p - (x + kL), dx^2 + dy^2 <= R^2) and records the hits as bits of a private bitmap.
//  phase 2   thread <-> grid point, integer only.  Each thread walks its set bits; for a particle i in the disc it
//            walks i's half-bonds i -> j.  A bond whose other end j is also in the disc (one bitmap probe) would be
//            visited twice by the reference, from i and from j, each time with weight 1/2 and the same line integral:
//            it is kept ONCE (from the end with the smaller local index) with weight 1; a bond with only i inside
//            keeps weight 1/2.  That is the reference's sum (IrvingKirkwood.h:80-107) with U(x) instead of ~1.5 U(x)
//            quadratures.  Survivors are appended to a flat item list of the tile (count, scan, emit), 4 bytes each:
//            (grid point, particle, bond, weight).
//  phase 3   thread <-> equal share of the item list.  Every lane integrates the same number of items, whatever the
//            local density: t(s) = c0 + s (c1 + s c2), Gauss-Legendre sum of the shape, 3 FMAs per item into register
//            accumulators; an item list is sorted by grid point, so a thread's share spans one or two grid points and
//            the partial sums are handed to the owning thread through shared memory in a fixed order.
//
// Everything discrete (disc membership, bond set) is bit-identical to the reference arithmetic; only the order of
// the floating-point sums differs.  A CTA whose tile does not meet the static limits below runs `ikt_slow_path`,
// the plain half-bond formulation straight from global memory -- same results, no limits.
#define IKT_TI 16
#define IKT_TJ 16
#define IKT_THREADS (IKT_TI * IKT_TJ)
#define IKT_MAXROWS 24
#define IKT_MAXRUNS (2 * IKT_MAXROWS)
#define IKT_MAXCOLS 40
#define IKT_NLOC 512
#define IKT_NW (IKT_NLOC / 32)
#define IKT_NOJ 0xFFFFu
#define IKT_MAXNB 64                     // half-bonds of one particle an item can address (6 bits)

struct IKTStatic {
    int ps[IKT_MAXRUNS];                 // first sorted particle of the run
    int np[IKT_MAXRUNS];                 // particles in the run
    int h0[IKT_MAXRUNS];                 // first half-bond record of the run (global)
    int lp_base[IKT_MAXRUNS + 1];        // tile-local index of the run's first particle
    int lr_base[IKT_MAXRUNS + 1];        // tile-local index of the run's first record
    double2 shift[IKT_MAXRUNS];          // (kx Lx, ky Ly) of the run's periodic image
    unsigned short lcs[IKT_MAXROWS][IKT_MAXCOLS + 2];   // local index of the first particle of cell (row, col)
    double2 pos[IKT_NLOC];               // stored (wrapped) positions; after phase 1: positions in the tile's image
    int lrec[IKT_NLOC + 1];              // tile-local record offset of each particle
    int hoff[IKT_NLOC];                  // global record offset (hb_off) of each particle
    unsigned char run_of[IKT_NLOC];
    unsigned bits[IKT_NW * IKT_THREADS]; // [word][thread]: particles inside the thread's disc
    double2 gxy[IKT_THREADS];            // grid point of each thread
    int off[IKT_THREADS + 1];            // first item of each grid point (exclusive scan of the item counts)
    int wsum[IKT_THREADS / 32];
    double mid[IKT_THREADS][3];          // sums of grid points that lie strictly inside one thread's item range
};

// last index r in [0, n) with base[r] <= v  (base is non-decreasing, base[0] <= v)
__device__ __forceinline__ int ikt_find_run(const int* base, int n, int v)
{
    int lo = 0, hi = n;
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (base[mid] <= v) lo = mid; else hi = mid;
    }
    return lo;
}

// The reference's own formulation for one grid point, from global memory (no tile limits): every half-bond of
// every particle in the disc, weight 1/2.
template <int KERNEL, int Q>
__device__ __forceinline__ void ikt_slow_path(const PgtIKArgs& A, double gx, double gy, double& oxx, double& oxy, double& oyy)
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
                    const double wf = bond_quadrature<KERNEL, Q>(c0, c1, cf.x, A.nq) * cf.y * 0.5;
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


// Gauss-Legendre sum of the shape along the bond.  The 5-point octic case is written out: nodes and weights come
// straight from the kernel-parameter constant bank, out-of-disc nodes are dropped by predication, two partial sums.
template <int KERNEL, int Q>
__device__ __forceinline__ double ikt_quadrature(const PgtIKArgs& A, double c0, double c1, double c2)
{
    if (KERNEL == PGT_CG_OCTIC && Q == 5) {
        double B0 = 0.0, B1 = 0.0;
#pragma unroll
        for (int k = 0; k < 5; ++k) {
            const double s = A.gl_s[k];
            const double t = fma(fma(c2, s, c1), s, c0);
            const double g = fma(-t, t, 1.0);
            const double gg = g * g;
            double& B = (k & 1) ? B1 : B0;
            asm("{\n\t.reg .pred p;\n\tsetp.lt.s32 p, %1, 1072693248;\n\t@p fma.rn.f64 %0, %2, %3, %0;\n\t}"
                : "+d"(B)
                : "r"(__double2hiint(t)), "d"(A.gl_w[k]), "d"(gg));
        }
        return B0 + B1;
    }
    return bond_quadrature<KERNEL, Q>(c0, c1, c2, A.nq);
}

template <int KERNEL, int Q>
__global__ void __launch_bounds__(IKT_THREADS, 3) k_ik_tile(const PgtIKArgs A, const int tile_i_begin, const int jl_cap,
                                                            const int item_cap)
{
    __shared__ IKTStatic S;
    extern __shared__ __align__(16) unsigned char ikt_dyn[];
    unsigned* const s_items = reinterpret_cast<unsigned*>(ikt_dyn);                    // item_cap words (also the hand-over area)
    unsigned short* const s_jl = reinterpret_cast<unsigned short*>(s_items + item_cap);  // tile-local index of each bond's j

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const PgtFrameView& fr = A.fr;
    const double R = A.cg.rcut, R2 = A.cg.rcutsq, iR2 = A.inv_rcutsq;
    const int ncx = A.cells.ncx, ncy = A.cells.ncy;

    // ---- tile and my grid point: a warp covers a compact 4 x 8 patch
    const int tiles_j = (A.grid.ny + IKT_TJ - 1) / IKT_TJ;
    const int i0 = (tile_i_begin + blockIdx.x / tiles_j) * IKT_TI, j0 = (blockIdx.x % tiles_j) * IKT_TJ;
    const int i1 = min(i0 + IKT_TI, A.grid.nx) - 1, j1 = min(j0 + IKT_TJ, A.grid.ny) - 1;
    const int gi = i0 + 4 * (warp >> 1) + (lane >> 3), gj = j0 + 8 * (warp & 1) + (lane & 7);
    // every in-grid point of the tile is computed, whatever the shard, so that a point's summation order (and
    // therefore its bits) does not depend on how the grid is split over GPUs; only in-shard points are stored
    const bool valid = gi <= i1 && gj <= j1;
    bool store = false;
    long long g = 0;
    double gx = 0.0, gy = 0.0;
    if (valid) {
        g = (long long)gi * A.grid.ny + gj;
        store = g >= A.grid.g0 && g < A.grid.g1;
        gx = __ldg(A.grid.xs + gi);
        gy = __ldg(A.grid.ys + gj);
    }
    S.gxy[tid] = make_double2(gx, gy);

    // ---- cell rectangle of the tile (unwrapped cell coordinates), same padding rule as a single point's window
    double xmin = __ldg(A.grid.xs + i0), xmax = xmin, ymin = __ldg(A.grid.ys + j0), ymax = ymin;
    for (int i = i0 + 1; i <= i1; ++i) { const double v = __ldg(A.grid.xs + i); xmin = fmin(xmin, v); xmax = fmax(xmax, v); }
    for (int j = j0 + 1; j <= j1; ++j) { const double v = __ldg(A.grid.ys + j); ymin = fmin(ymin, v); ymax = fmax(ymax, v); }
    const double tpx = 1e-9 * (R + fmax(fabs(xmin), fabs(xmax)) + 1.0), tpy = 1e-9 * (R + fmax(fabs(ymin), fabs(ymax)) + 1.0);
    const int ux0 = __double2int_rd((xmin - R - tpx - A.box.lo[0]) * A.cells.inv_cwx);
    const int ux1 = __double2int_rd((xmax + R + tpx - A.box.lo[0]) * A.cells.inv_cwx);
    const int uy0 = __double2int_rd((ymin - R - tpy - A.box.lo[1]) * A.cells.inv_cwy);
    const int uy1 = __double2int_rd((ymax + R + tpy - A.box.lo[1]) * A.cells.inv_cwy);
    const int ncols = ux1 - ux0 + 1, nrows = uy1 - uy0 + 1;
    // static limits; the rectangle plus a bond's reach must not wrap onto itself (one image of every particle)
    bool fast = ncols <= IKT_MAXCOLS && nrows <= IKT_MAXROWS && ncols + 2 * A.cells.reach + 1 <= ncx &&
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
        if (warp == 0) {                                   // exclusive scans over <= 48 runs: lane <-> row
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
        if (nloc > IKT_NLOC || nrec_total > jl_cap) fast = false;     // uniform: every thread reads the same totals
    }

    int toobig = 0;
    if (fast) {
        // ---- local cell starts
        for (int e = tid; e < nrows * (ncols + 1); e += IKT_THREADS) {
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
            S.lcs[r][c] = (unsigned short)v;
        }
        // ---- particles
        for (int l = tid; l < nloc; l += IKT_THREADS) {
            const int run = ikt_find_run(S.lp_base, nruns, l);
            const int p = S.ps[run] + (l - S.lp_base[run]);
            S.pos[l] = make_double2(__ldg(fr.sx + p), __ldg(fr.sy + p));
            const int h = __ldg(fr.hb_off + p), hn = __ldg(fr.hb_off + p + 1);
            S.lrec[l] = S.lr_base[run] + (h - S.h0[run]);
            S.hoff[l] = h;
            S.run_of[l] = (unsigned char)run;
            if (hn - h > IKT_MAXNB) toobig = 1;
        }
        if (tid == 0) S.lrec[nloc] = nrec_total;
        // ---- tile-local index of every bond's other end (IKT_NOJ when it lies outside the rectangle)
        for (int k = tid; k < nrec_total; k += IKT_THREADS) {
            const int run = ikt_find_run(S.lr_base, nruns, k);
            const int j = __ldg(fr.hb_j + S.h0[run] + (k - S.lr_base[run]));
            // j lies within `reach` cell rows of i: probe those rows' runs for the one that holds it
            const int r = run >> 1;
            unsigned jl = IKT_NOJ;
            const int q0 = 2 * max(r - A.cells.reach, 0), q1 = 2 * min(r + A.cells.reach, nrows - 1) + 1;
            for (int q = q0; q <= q1; ++q) {
                const unsigned o = (unsigned)(j - S.ps[q]);
                if (o < (unsigned)S.np[q]) { jl = (unsigned)S.lp_base[q] + o; break; }
            }
            s_jl[k] = (unsigned short)jl;
        }
#pragma unroll
        for (int w = 0; w < IKT_NW; ++w) S.bits[w * IKT_THREADS + tid] = 0u;
    }
    if (__syncthreads_or(toobig)) fast = false;            // a particle with more bonds than an item can address

    double txx = 0.0, txy = 0.0, tyy = 0.0;
    const double out_scale = A.cg.norm;                    // weights 1 / 0.5 already carry VirialStress::compute_cg's 1/2
    double* const out_ptr = A.out_virial + 4 * (g - A.grid.g0);

    // ================= phase 1: my disc as a bitmap over tile-local particle indices =================
    int w_first = IKT_NW, w_last = -1;
    if (fast && valid) {
        const double padx = 1e-9 * (R + fabs(gx) + 1.0), pady = 1e-9 * (R + fabs(gy) + 1.0);
        const int cx0 = max(__double2int_rd((gx - R - padx - A.box.lo[0]) * A.cells.inv_cwx), ux0) - ux0;
        const int cx1 = min(__double2int_rd((gx + R + padx - A.box.lo[0]) * A.cells.inv_cwx), ux1) - ux0;
        const int ry0 = max(__double2int_rd((gy - R - pady - A.box.lo[1]) * A.cells.inv_cwy), uy0) - uy0;
        const int ry1 = min(__double2int_rd((gy + R + pady - A.box.lo[1]) * A.cells.inv_cwy), uy1) - uy0;
        int cur_w = -1;
        unsigned cur_bits = 0u;
        for (int r = ry0; r <= ry1; ++r) {
            const int la = S.lcs[r][cx0], lb = S.lcs[r][cx1 + 1], lsplit = S.lp_base[2 * r + 1];
            const double2 shA = S.shift[2 * r], shB = S.shift[2 * r + 1];
            const double cpy = gy + shA.y, cpxA = gx + shA.x, cpxB = gx + shB.x;     // x + k (bmax - bmin), Search.h:186-189
            for (int l = la; l < lb; ++l) {
                const double2 p = S.pos[l];
                const double ax = p.x - (l < lsplit ? cpxA : cpxB);
                const double ay = p.y - cpy;
                const double d2 = ax * ax + ay * ay;
                if (!(d2 > R2)) {
                    const int w = l >> 5;
                    if (w != cur_w) {
                        if (cur_w >= 0) S.bits[cur_w * IKT_THREADS + tid] = cur_bits;
                        else w_first = w;
                        cur_w = w;
                        cur_bits = 0u;
                    }
                    cur_bits |= 1u << (l & 31);
                }
            }
        }
        if (cur_w >= 0) { S.bits[cur_w * IKT_THREADS + tid] = cur_bits; w_last = cur_w; }
    }

    // ================= phase 2: which (particle, bond) items are mine =================
    // walk(f) calls f(l, r, both) for every surviving item of my grid point, in ascending (l, r)
    const unsigned* const mybits = S.bits + tid;
    auto walk = [&](auto&& f) {
        for (int w = w_first; w <= w_last; ++w) {
            unsigned bits = mybits[w * IKT_THREADS];
            while (bits) {
                const int l = (w << 5) + __ffs(bits) - 1;
                bits &= bits - 1u;
                const int k0 = S.lrec[l], k1 = S.lrec[l + 1];
                for (int k = k0; k < k1; ++k) {
                    const unsigned jl = s_jl[k];
                    const bool both = jl != IKT_NOJ && ((mybits[(jl >> 5) * IKT_THREADS] >> (jl & 31)) & 1u);
                    if (both && jl < (unsigned)l) continue;         // kept from j's side
                    f(l, k - k0, both);
                }
            }
        }
    };
    int n_mine = 0;
    if (fast) {
        __syncthreads();                                   // s_jl complete (bitmaps are private)
        walk([&](int, int, bool) { ++n_mine; });
    }
    if (__syncthreads_or(n_mine > item_cap)) fast = false; // also orders phase 1's reads of S.pos before the rewrite below

    if (!fast) {
        if (store) {
            ikt_slow_path<KERNEL, Q>(A, gx, gy, txx, txy, tyy);
            *reinterpret_cast<double4*>(out_ptr) = make_double4(out_scale * txx, out_scale * txy, out_scale * txy, out_scale * tyy);
        }
        return;
    }

    // exclusive scan of the item counts over the 256 grid points of the tile
    {
        int v = n_mine;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int t = __shfl_up_sync(FULL_MASK, v, d);
            if (lane >= d) v += t;
        }
        if (lane == 31) S.wsum[warp] = v;
        // positions of the tile's periodic image: r - kL (used for the quadrature only, never for a membership test)
        for (int l = tid; l < nloc; l += IKT_THREADS) {
            const double2 p = S.pos[l], sh = S.shift[S.run_of[l]];
            S.pos[l] = make_double2(p.x - sh.x, p.y - sh.y);
        }
        __syncthreads();
        int base = 0;
#pragma unroll
        for (int w = 0; w < IKT_THREADS / 32; ++w) base += (w < warp) ? S.wsum[w] : 0;
        S.off[tid] = base + v - n_mine;
        if (tid == IKT_THREADS - 1) S.off[IKT_THREADS] = base + v;
        __syncthreads();
    }
    const int my_off = S.off[tid], n_items = S.off[IKT_THREADS];

    // ================= phase 3: equal shares of the item list =================
    double* const hand = reinterpret_cast<double*>(s_items);       // [thread][6]: head and tail partial sums
    for (int P0 = 0; P0 < IKT_THREADS;) {
        // grid points [P0, P1) whose items fit the buffer (uniform)
        const int ibase = S.off[P0];
        int P1;
        if (n_items - ibase <= item_cap) P1 = IKT_THREADS;
        else {
            int lo = P0 + 1, hi = IKT_THREADS;             // off[lo] - ibase <= cap holds: one point's items fit (checked above)
            while (hi > lo) {
                const int mid = (lo + hi + 1) >> 1;
                if (S.off[mid] - ibase <= item_cap) lo = mid; else hi = mid - 1;
            }
            P1 = lo;
        }
        const int m = S.off[P1] - ibase;
        if (tid >= P0 && tid < P1) {
            unsigned* dst = s_items + (my_off - ibase);
            walk([&](int l, int r, bool both) { *dst++ = (unsigned)tid | ((unsigned)l << 8) | ((unsigned)r << 17) | ((both ? 1u : 0u) << 23); });
        }
        __syncthreads();

        // my share: items [tid q, tid q + q); q odd -> the 32 lanes of a warp read 32 different banks
        const int q = ((m + IKT_THREADS - 1) / IKT_THREADS) | 1;
        const int ia = tid * q, ib = min(ia + q, m);
        double axx = 0.0, axy = 0.0, ayy = 0.0, hxx = 0.0, hxy = 0.0, hyy = 0.0;
        int cur_p = -1, nflush = 0;
        for (int it = ia; it < ib; ++it) {
            const unsigned item = s_items[it];
            const int p = item & 255u, l = (item >> 8) & 511u, r = (item >> 17) & 63u;
            if (p != cur_p) {
                if (cur_p >= 0) {
                    if (nflush == 0) { hxx = axx; hxy = axy; hyy = ayy; }
                    else { S.mid[cur_p][0] = axx; S.mid[cur_p][1] = axy; S.mid[cur_p][2] = ayy; }
                    ++nflush;
                }
                cur_p = p;
                axx = axy = ayy = 0.0;
            }
            const double2 gp = S.gxy[p], u = S.pos[l];
            const int e = S.hoff[l] + r;
            const double2 b = __ldg(fr.hb_b + e), cf = __ldg(fr.hb_cf + e);
            const double ax = u.x - gp.x, ay = u.y - gp.y;
            const double c0 = fma(ax, ax, ay * ay) * iR2;
            const double c1 = fma(ax, b.x, ay * b.y) * (2.0 * iR2);
            const double wgt = (item & (1u << 23)) ? 1.0 : 0.5;
            const double wf = ikt_quadrature<KERNEL, Q>(A, c0, c1, cf.x) * (cf.y * wgt);
            const double tx = wf * b.x, ty = wf * b.y;
            axx = fma(tx, b.x, axx);
            axy = fma(tx, b.y, axy);
            ayy = fma(ty, b.y, ayy);
        }
        if (nflush == 0) { hxx = axx; hxy = axy; hyy = ayy; }     // my only grid point: head = tail
        __syncthreads();                                   // every share read: the buffer becomes the hand-over area
        hand[tid * 6 + 0] = hxx; hand[tid * 6 + 1] = hxy; hand[tid * 6 + 2] = hyy;
        hand[tid * 6 + 3] = axx; hand[tid * 6 + 4] = axy; hand[tid * 6 + 5] = ayy;
        __syncthreads();
        // owner of grid point `tid`: add the partial sums of the threads that covered its items, in thread order
        if (tid >= P0 && tid < P1 && n_mine > 0) {
            const int a = my_off - ibase, bnd = a + n_mine;
            const int ta = a / q, tb = (bnd - 1) / q;
            for (int t = ta; t <= tb; ++t) {
                const int sa = t * q, sb = min(sa + q, m);
                const double* src;
                if (a <= sa) src = hand + t * 6;                  // first grid point of thread t's share
                else if (bnd >= sb) src = hand + t * 6 + 3;       // last one
                else src = &S.mid[tid][0];                        // strictly inside
                txx += src[0]; txy += src[1]; tyy += src[2];
            }
        }
        __syncthreads();
        P0 = P1;
    }

    if (store) *reinterpret_cast<double4*>(out_ptr) = make_double4(out_scale * txx, out_scale * txy, out_scale * txy, out_scale * tyy);
}

static bool ik_tile_applicable(const PgtIKArgs& a)
{
    return a.grid.mode == 0 && a.obs_mask == PGT_OBS_VIRIAL && a.out_virial != nullptr;
}

static int g_ikt_item_cap = 0, g_ikt_jl_cap = 0;    // sized once from the device's shared memory

template <int KERNEL, int Q>
static cudaError_t launch_ik_tile_t(const PgtIKArgs& a, cudaStream_t st)
{
    if (a.grid.g1 <= a.grid.g0) return cudaSuccess;
    if (g_ikt_item_cap == 0) {
        int dev = 0, smem_sm = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&smem_sm, cudaDevAttrMaxSharedMemoryPerMultiprocessor, dev);
        // three CTAs per SM: a third of the SM's shared memory minus the static part and the per-CTA reservation
        const long long budget = (long long)smem_sm / 3 - (long long)sizeof(IKTStatic) - 1280;
        g_ikt_jl_cap = 4096;
        g_ikt_item_cap = (int)std::max<long long>(1536, (budget - 2LL * g_ikt_jl_cap) / 4) & ~3;
    }
    const size_t dyn = (size_t)g_ikt_item_cap * 4 + (size_t)g_ikt_jl_cap * 2;
    static bool attr_set_dev[64] = { false };               // one flag per template instance and device
    int cur_dev = 0;
    cudaGetDevice(&cur_dev);
    bool& attr_set = attr_set_dev[cur_dev & 63];
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(k_ik_tile<KERNEL, Q>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn);
        if (e != cudaSuccess) return e;
        attr_set = true;
    }
    const int tiles_j = (a.grid.ny + IKT_TJ - 1) / IKT_TJ;
    const int i_lo = (int)(a.grid.g0 / a.grid.ny), i_hi = (int)((a.grid.g1 - 1) / a.grid.ny);
    const int tile_i_begin = i_lo / IKT_TI;
    const long long blocks = (long long)(i_hi / IKT_TI - tile_i_begin + 1) * tiles_j;
    k_ik_tile<KERNEL, Q><<<(unsigned)blocks, IKT_THREADS, dyn, st>>>(a, tile_i_begin, g_ikt_jl_cap, g_ikt_item_cap);
    return cudaGetLastError();
}

template <int KERNEL>
static cudaError_t launch_ik_tile_q(const PgtIKArgs& a, cudaStream_t st)
{
    if (a.nq == 5) return launch_ik_tile_t<KERNEL, 5>(a, st);
    return launch_ik_tile_t<KERNEL, 0>(a, st);
}

static cudaError_t launch_ik_tile(const PgtIKArgs& a, cudaStream_t st)
{
    switch (a.cg.kernel) {
    case PGT_CG_OCTIC: return launch_ik_tile_q<PGT_CG_OCTIC>(a, st);
    case PGT_CG_RECT: return launch_ik_tile_q<PGT_CG_RECT>(a, st);
    default: return launch_ik_tile_q<PGT_CG_MOLLIFIER>(a, st);
    }
}
