// pgt_ik_pairtile.inl -- stage 4/5, third formulation ("pair tile"); included by pgt_kernels.cu.
//
// The work of the gather has two levels: (grid point x, particle p in the disc of x) pairs, ~13 per
// grid point, each with its own a = r_p - x, and under every pair the n_b(p) half-bonds of p.  The
// earlier kernels paid a per-bond lookup of the pair data; here a LANE OWNS A PAIR and loops over its
// half-bonds, so the inner loop is only record loads + FP64:
//
//  phase 1  warp <-> patch of 4 x 8 grid points, lane <-> grid point.  The warp walks the cell rows of
//           the patch's bounding window once; every candidate particle is loaded warp-uniformly and
//           tested by all 32 lanes exactly as Aboria does (dx = r_p - (x + kL), dx^2+dy^2 <= R^2,
//           Search.h:186-193); one ballot gives the mask of grid points that see it.  Particles with a
//           non-empty mask become entries (position, image code, first half-bond, n_b, mask) in shared
//           memory.  ~12 instructions per candidate for 32 tests.
//  sort     the entries of a batch are counting-sorted by n_b (descending, stable) with match_any, and
//           the set bits of their masks are expanded into a pair list: pairs with equal trip count and
//           equal particle (= equal record addresses, broadcast loads) are adjacent.
//  phase 2  lane <-> pair, 32 pairs per round: a, |a|^2/R^2, 2a/R^2 once per pair, then n_b iterations of
//           {3 x 16-byte record loads, 27 FP64 instructions}: t(s) at the 5 nodes in symmetric form
//           A +- h B + h^2 c2, shape (1-t^2)^2 under a sign-word test, three FMAs into the pair's sums.
//  deliver  the pair sums are added into the per-grid-point accumulators of the warp in shared memory,
//           conflicting lanes (same grid point) serialised in lane order: the summation order for a grid
//           point is (batch, n_b descending, scan order of p), the same for every sharding of the grid.
//  store    lane <-> grid point: one 32-byte store per point, 8 consecutive points per row segment.
#define IK3_WARPS 4
#define IK3_THREADS (IK3_WARPS * 32)
#define IK3_PI 4                       // patch rows   (i, slow index)
#define IK3_PJ 8                       // patch columns (j, fast index)
#define IK3_TILE_I (2 * IK3_PI)        // 2 x 2 patches per CTA: neighbouring patches share L1 lines
#define IK3_TILE_J (2 * IK3_PJ)
#define IK3_ECAP 96                    // entries per batch
#define IK3_PCAP 1536                  // pairs per batch
#define IK3_NBINS 32

struct IK3WarpSmem {
    double2 ent_xy[IK3_ECAP];          // stored (wrapped) position
    int ent_hb0[IK3_ECAP];             // first half-bond
    int ent_q[IK3_ECAP];               // sorted particle index (local observables)
    unsigned ent_mask[IK3_ECAP];       // grid points of the patch that have it in their disc
    unsigned short ent_nb[IK3_ECAP];   // half-bonds
    unsigned short pair_start[IK3_ECAP];
    unsigned char ent_k[IK3_ECAP];     // image code (kx+1) | (ky+1) << 2
    unsigned char perm[IK3_ECAP];      // sorted position -> entry
    unsigned char rankin[IK3_ECAP];    // rank of the entry inside its bin
    unsigned short pairs[IK3_PCAP];    // (sorted position << 5) | lane of the grid point
    int bin[IK3_NBINS];
    double acc[7][32];                 // xx, xy, yy | kxx, kxy, kyy, rho
    double gx[32], gy[32];
};

__device__ __forceinline__ int ik3_bin(int nb)
{
    // descending n_b <-> ascending key; long lists share coarse bins (their rounds run to the round's max)
    const int k = nb < 24 ? nb : min(IK3_NBINS - 1, 24 + ((nb - 24) >> 4));
    return IK3_NBINS - 1 - k;
}

// 5-point Gauss-Legendre sum of the octic shape along the bond, symmetric form.
__device__ __forceinline__ double bond_quad_octic5(const PgtIKArgs& A, double c0, double c1, double c2)
{
    const double Am = fma(0.25, c2, fma(0.5, c1, c0));      // t(1/2)
    const double Bq = c1 + c2;                              // t'(1/2)
    const double g0 = fma(-Am, Am, 1.0);
    double B = 0.0;
    if (__double2hiint(Am) < 0x3FF00000) B = A.gl5_w0 * (g0 * g0);
    {
        const double m = fma(A.gl5_h1sq, c2, Am);
        const double tp = fma(A.gl5_h1, Bq, m), tm = fma(-A.gl5_h1, Bq, m);
        const double gp = fma(-tp, tp, 1.0), gm = fma(-tm, tm, 1.0);
        double q = 0.0;
        if (__double2hiint(tp) < 0x3FF00000) q = gp * gp;
        if (__double2hiint(tm) < 0x3FF00000) q = fma(gm, gm, q);
        B = fma(A.gl5_w1, q, B);
    }
    {
        const double m = fma(A.gl5_h2sq, c2, Am);
        const double tp = fma(A.gl5_h2, Bq, m), tm = fma(-A.gl5_h2, Bq, m);
        const double gp = fma(-tp, tp, 1.0), gm = fma(-tm, tm, 1.0);
        double q = 0.0;
        if (__double2hiint(tp) < 0x3FF00000) q = gp * gp;
        if (__double2hiint(tm) < 0x3FF00000) q = fma(gm, gm, q);
        B = fma(A.gl5_w2, q, B);
    }
    return B;
}

// Sort the batch, expand it into pairs, run the pairs, deliver into sm.acc.
template <int KERNEL, int Q, bool LOCAL>
__device__ __forceinline__ void ik3_flush(const PgtIKArgs& A, IK3WarpSmem& sm, const int lane, const int ne, const int npairs)
{
    const PgtFrameView& fr = A.fr;
    const unsigned lt_mask = (1u << lane) - 1u;
    const double iR2 = A.inv_rcutsq, two_iR2 = 2.0 * A.inv_rcutsq;
    const double Lx = A.box.L[0], Ly = A.box.L[1];

    // ---- stable counting sort of the entries by bin(n_b)
    sm.bin[lane] = 0;
    __syncwarp();
    for (int e0 = 0; e0 < ne; e0 += 32) {
        const int e = e0 + lane;
        const bool act = e < ne;
        const int key = act ? ik3_bin(sm.ent_nb[e]) : IK3_NBINS + lane;
        const unsigned peers = __match_any_sync(FULL_MASK, key);
        const int rank = __popc(peers & lt_mask);
        int old = 0;
        if (act && rank == 0) {                      // one leader per bin present in this round
            old = sm.bin[key];
            sm.bin[key] = old + __popc(peers);
        }
        old = __shfl_sync(FULL_MASK, old, __ffs(peers) - 1);
        if (act) sm.rankin[e] = (unsigned char)(old + rank);
        __syncwarp();
    }
    {
        const int cnt = sm.bin[lane];
        int inc = cnt;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int t = __shfl_up_sync(FULL_MASK, inc, d);
            if (lane >= d) inc += t;
        }
        __syncwarp();
        sm.bin[lane] = inc - cnt;
        __syncwarp();
    }
    for (int e = lane; e < ne; e += 32) sm.perm[sm.bin[ik3_bin(sm.ent_nb[e])] + sm.rankin[e]] = (unsigned char)e;
    __syncwarp();

    // ---- pair offsets in sorted order, then expansion of the masks
    {
        int running = 0;
        for (int s0 = 0; s0 < ne; s0 += 32) {
            const int s = s0 + lane;
            const int c = s < ne ? __popc(sm.ent_mask[sm.perm[s]]) : 0;
            int inc = c;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int t = __shfl_up_sync(FULL_MASK, inc, d);
                if (lane >= d) inc += t;
            }
            if (s < ne) sm.pair_start[s] = (unsigned short)(running + inc - c);
            running += __shfl_sync(FULL_MASK, inc, 31);
        }
    }
    __syncwarp();
    for (int s = lane; s < ne; s += 32) {
        unsigned m = sm.ent_mask[sm.perm[s]];
        int o = sm.pair_start[s];
        while (m) {
            const int b = __ffs(m) - 1;
            m &= m - 1;
            sm.pairs[o++] = (unsigned short)((s << 5) | b);
        }
    }
    __syncwarp();

    // ---- phase 2: lane <-> pair
    for (int t0 = 0; t0 < npairs; t0 += 32) {
        const int t = t0 + lane;
        const bool act = t < npairs;
        int xl = 0, nb = 0, hb0 = 0;
        double c0 = 0.0, a2x = 0.0, a2y = 0.0;
        double lrho = 0.0, lkxx = 0.0, lkxy = 0.0, lkyy = 0.0;
        if (act) {
            const unsigned pr = sm.pairs[t];
            xl = pr & 31;
            const int e = sm.perm[pr >> 5];
            const double2 pxy = sm.ent_xy[e];
            const int kc = sm.ent_k[e];
            hb0 = sm.ent_hb0[e];
            nb = sm.ent_nb[e];
            const int kx = (kc & 3) - 1, ky = (kc >> 2) - 1;
            const double ax = pxy.x - (sm.gx[xl] + kx * Lx);       // as Aboria forms it
            const double ay = pxy.y - (sm.gy[xl] + ky * Ly);
            const double d2 = ax * ax + ay * ay;
            c0 = d2 * iR2;
            a2x = ax * two_iR2;
            a2y = ay * two_iR2;
            if (LOCAL) {
                const double cgval = pgt_cg_delta_exact(A.cg, d2);
                lrho = cgval;
                if (fr.skxx) {
                    const int q = sm.ent_q[e];
                    lkxx = __ldg(fr.skxx + q) * cgval;
                    lkxy = __ldg(fr.skxy + q) * cgval;
                    lkyy = __ldg(fr.skyy + q) * cgval;
                }
            }
        }
        const int nbmax = __reduce_max_sync(FULL_MASK, nb);
        double pxx = 0.0, pxy_ = 0.0, pyy = 0.0;
        const double2* rec = fr.hb_rec + (size_t)hb0 * PGT_REC_STRIDE;
        for (int k = 0; k < nbmax; ++k) {
            if (k < nb) {
                const double2 r0 = __ldg(rec + k * PGT_REC_STRIDE);
                const double2 r1 = __ldg(rec + k * PGT_REC_STRIDE + 1);
                const double2 r2 = __ldg(rec + k * PGT_REC_STRIDE + 2);
                const double c1 = fma(a2x, r0.x, a2y * r0.y);
                double B;
                if (KERNEL == PGT_CG_OCTIC && Q == 5) B = bond_quad_octic5(A, c0, c1, r1.x);
                else B = bond_quadrature<KERNEL, Q>(c0, c1, r1.x, A.nq);
                pxx = fma(B, r1.y, pxx);
                pxy_ = fma(B, r2.x, pxy_);
                pyy = fma(B, r2.y, pyy);
            }
        }

        // ---- deliver: lanes that hold the same grid point add in lane order
        const unsigned peers = __match_any_sync(FULL_MASK, act ? xl : 32 + lane);
        const int rank = __popc(peers & lt_mask);
        const int maxm = __reduce_max_sync(FULL_MASK, act ? __popc(peers) : 0);
        for (int r = 0; r < maxm; ++r) {
            if (act && rank == r) {
                sm.acc[0][xl] += pxx;
                sm.acc[1][xl] += pxy_;
                sm.acc[2][xl] += pyy;
                if (LOCAL) {
                    sm.acc[3][xl] += lkxx;
                    sm.acc[4][xl] += lkxy;
                    sm.acc[5][xl] += lkyy;
                    sm.acc[6][xl] += lrho;
                }
            }
            __syncwarp();
        }
    }
    __syncwarp();      // the entry tables may be overwritten by the next batch from here on
}

template <int KERNEL, int Q, bool LOCAL>
__global__ void __launch_bounds__(IK3_THREADS, 4) k_ik_gather3(const PgtIKArgs A, const int tile_i_begin)
{
    __shared__ IK3WarpSmem smem_all[IK3_WARPS];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    IK3WarpSmem& sm = smem_all[warp];
    const PgtFrameView& fr = A.fr;
    const double R = A.cg.rcut, R2 = A.cg.rcutsq;
    const double Lx = A.box.L[0], Ly = A.box.L[1];
    const int ncx = A.cells.ncx, ncy = A.cells.ncy;
    const bool want_virial = (A.obs_mask & PGT_OBS_VIRIAL) != 0;

    // ---- my grid point.  in_grid lanes take part in the search whatever the shard, so that the entry
    //      batches (and with them every summation order) do not depend on how the grid is sharded.
    const int tiles_j = (A.grid.ny + IK3_TILE_J - 1) / IK3_TILE_J;
    const int ti = tile_i_begin + blockIdx.x / tiles_j, tj = blockIdx.x % tiles_j;
    const int i = ti * IK3_TILE_I + (warp >> 1) * IK3_PI + (lane >> 3);
    const int j = tj * IK3_TILE_J + (warp & 1) * IK3_PJ + (lane & 7);
    const bool in_grid = i < A.grid.nx && j < A.grid.ny;
    const long long g = (long long)i * A.grid.ny + j;
    const bool valid = in_grid && g >= A.grid.g0 && g < A.grid.g1;
    if (!__any_sync(FULL_MASK, valid)) return;
    double gx = 0.0, gy = 0.0;
    if (in_grid) { gx = __ldg(A.grid.xs + i); gy = __ldg(A.grid.ys + j); }
    sm.gx[lane] = gx;
    sm.gy[lane] = gy;
#pragma unroll
    for (int k = 0; k < 7; ++k) sm.acc[k][lane] = 0.0;

    // ---- the patch's cell window: union of the lanes' windows (unwrapped cell coordinates)
    int ux0 = INT_MAX, ux1 = INT_MIN, uy0 = INT_MAX, uy1 = INT_MIN;
    if (in_grid) {
        const double padx = 1e-9 * (R + fabs(gx) + 1.0), pady = 1e-9 * (R + fabs(gy) + 1.0);
        ux0 = max(__double2int_rd((gx - R - padx - A.box.lo[0]) * A.cells.inv_cwx), -ncx);
        ux1 = min(__double2int_rd((gx + R + padx - A.box.lo[0]) * A.cells.inv_cwx), 2 * ncx - 1);
        uy0 = max(__double2int_rd((gy - R - pady - A.box.lo[1]) * A.cells.inv_cwy), -ncy);
        uy1 = min(__double2int_rd((gy + R + pady - A.box.lo[1]) * A.cells.inv_cwy), 2 * ncy - 1);
    }
    ux0 = __reduce_min_sync(FULL_MASK, ux0);
    ux1 = __reduce_max_sync(FULL_MASK, ux1);
    uy0 = __reduce_min_sync(FULL_MASK, uy0);
    uy1 = __reduce_max_sync(FULL_MASK, uy1);
    Seg seg[3];
    const int nseg = split_segments(ux0, ux1, ncx, seg);
    __syncwarp();

    // ---- phase 1: warp-uniform walk over the candidates, 32 exact tests per candidate
    int ne = 0, npairs = 0;
    for (int uy = uy0; uy <= uy1; ++uy) {
        int ky = 0, row = uy;
        if (uy < 0) { ky = +1; row += ncy; }
        else if (uy >= ncy) { ky = -1; row -= ncy; }
        const double cpy = gy + ky * Ly;                      // x + k (bmax - bmin), Search.h:186-189
        const int rbase = row * ncx;
        for (int s = 0; s < nseg; ++s) {
            Seg sg = seg[0];                                   // select, not index: keeps seg[] in registers
            if (s == 1) sg = seg[1];
            if (s == 2) sg = seg[2];
            const double cpx = gx + sg.k * Lx;
            const int qb = __ldg(fr.cell_start + rbase + sg.c0);
            const int qe = __ldg(fr.cell_start + rbase + sg.c1 + 1);
            const int kcode = (sg.k + 1) | ((ky + 1) << 2);
            double px = 0.0, py = 0.0;
            if (qb < qe) { px = __ldg(fr.sx + qb); py = __ldg(fr.sy + qb); }
            for (int q = qb; q < qe; ++q) {
                const double cx = px, cy = py;
                if (q + 1 < qe) { px = __ldg(fr.sx + q + 1); py = __ldg(fr.sy + q + 1); }   // next candidate in flight
                const double dx = cx - cpx, dy = cy - cpy;
                const double d2 = dx * dx + dy * dy;
                const unsigned m = __ballot_sync(FULL_MASK, in_grid && !(d2 > R2));
                if (m == 0) continue;
                int hb0 = 0, nb = 0;
                if (want_virial) {
                    hb0 = __ldg(fr.hb_off + q);
                    nb = min(__ldg(fr.hb_off + q + 1) - hb0, 65535);
                }
                if (!LOCAL && nb == 0) continue;
                if (lane == 0) {
                    sm.ent_xy[ne] = make_double2(cx, cy);
                    sm.ent_hb0[ne] = hb0;
                    sm.ent_q[ne] = q;
                    sm.ent_mask[ne] = m;
                    sm.ent_nb[ne] = (unsigned short)nb;
                    sm.ent_k[ne] = (unsigned char)kcode;
                }
                ++ne;
                npairs += __popc(m);
                if (ne == IK3_ECAP || npairs > IK3_PCAP - 32) {
                    __syncwarp();
                    ik3_flush<KERNEL, Q, LOCAL>(A, sm, lane, ne, npairs);
                    ne = 0;
                    npairs = 0;
                }
            }
        }
    }
    if (ne > 0) {
        __syncwarp();
        ik3_flush<KERNEL, Q, LOCAL>(A, sm, lane, ne, npairs);
    }
    __syncwarp();

    // ---- store: lane <-> grid point
    if (valid) {
        const long long lg = g - A.grid.g0;
        if (A.out_virial && want_virial) {
            const double scale = 0.5 * A.cg.norm;               // val += 0.5 * bondval * (f rij (x) rij)
            const double xx = sm.acc[0][lane], xy = sm.acc[1][lane], yy = sm.acc[2][lane];
            *reinterpret_cast<double4*>(A.out_virial + 4 * lg) = make_double4(scale * xx, scale * xy, scale * xy, scale * yy);
        }
        if (LOCAL && A.out_kinetic && (A.obs_mask & PGT_OBS_KINETIC))
            *reinterpret_cast<double4*>(A.out_kinetic + 4 * lg) =
                make_double4(sm.acc[3][lane], sm.acc[4][lane], sm.acc[4][lane], sm.acc[5][lane]);
        if (LOCAL && A.out_density && (A.obs_mask & PGT_OBS_DENSITY)) A.out_density[lg] = sm.acc[6][lane];
    }
}

template <int KERNEL, int Q>
static cudaError_t launch_ik3_t(const PgtIKArgs& a, cudaStream_t st)
{
    if (a.grid.g1 <= a.grid.g0) return cudaSuccess;
    const int tiles_j = (a.grid.ny + IK3_TILE_J - 1) / IK3_TILE_J;
    const int i_lo = (int)(a.grid.g0 / a.grid.ny), i_hi = (int)((a.grid.g1 - 1) / a.grid.ny);
    const int tile_i_begin = i_lo / IK3_TILE_I;
    const long long blocks = (long long)(i_hi / IK3_TILE_I - tile_i_begin + 1) * tiles_j;
    if (blocks <= 0) return cudaSuccess;
    const bool local = (a.obs_mask & (PGT_OBS_KINETIC | PGT_OBS_DENSITY)) != 0;
    if (local) k_ik_gather3<KERNEL, Q, true><<<(unsigned)blocks, IK3_THREADS, 0, st>>>(a, tile_i_begin);
    else k_ik_gather3<KERNEL, Q, false><<<(unsigned)blocks, IK3_THREADS, 0, st>>>(a, tile_i_begin);
    return cudaGetLastError();
}

template <int KERNEL>
static cudaError_t launch_ik3_q(const PgtIKArgs& a, cudaStream_t st)
{
    if (a.nq == 5) return launch_ik3_t<KERNEL, 5>(a, st);
    return launch_ik3_t<KERNEL, 0>(a, st);
}

static cudaError_t launch_ik3(const PgtIKArgs& a, cudaStream_t st)
{
    switch (a.cg.kernel) {
    case PGT_CG_OCTIC: return launch_ik3_q<PGT_CG_OCTIC>(a, st);
    case PGT_CG_RECT: return launch_ik3_q<PGT_CG_RECT>(a, st);
    default: return launch_ik3_q<PGT_CG_MOLLIFIER>(a, st);
    }
}
