// pgt_capi.cu -- context management and the C-ABI of libpgt_cuda.so (include/pgt_cuda.h).
#include "pgt_cuda.h"
#include "pgt_kernels.cuh"
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <pthread.h>
#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

namespace {

struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    template <class T> T* as() const { return reinterpret_cast<T*>(p); }
};

thread_local std::string g_create_error;

// --- NCCL, resolved lazily with dlopen so the library loads (and single-GPU runs work) without it
typedef struct { char internal[128]; } pgtNcclUniqueId;
typedef void* pgtNcclComm;
struct NcclApi {
    void* handle = nullptr;
    int (*GetUniqueId)(pgtNcclUniqueId*) = nullptr;
    int (*CommInitRank)(pgtNcclComm*, int, pgtNcclUniqueId, int) = nullptr;
    int (*CommDestroy)(pgtNcclComm) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int, int, pgtNcclComm, cudaStream_t) = nullptr;
    int (*Broadcast)(const void*, void*, size_t, int, int, pgtNcclComm, cudaStream_t) = nullptr;
    int (*AllGather)(const void*, void*, size_t, int, pgtNcclComm, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    bool ok = false;
};
NcclApi g_nccl;

bool load_nccl(std::string& err)
{
    if (g_nccl.ok) return true;
    // A copy the process has already loaded (a host program that bundles NCCL, e.g. torch) wins: two libnccl.so.2 in one
    // process break whichever is loaded second.  Otherwise the system copy by absolute path, then the loader's search.
    g_nccl.handle = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
    const char* names[] = { "/usr/lib/x86_64-linux-gnu/libnccl.so.2", "libnccl.so.2", "libnccl.so" };
    for (const char* n : names) {
        if (g_nccl.handle) break;
        g_nccl.handle = dlopen(n, RTLD_NOW | RTLD_LOCAL);    // local: never interposes on a copy loaded later
    }
    if (!g_nccl.handle) { err = std::string("cannot dlopen libnccl: ") + dlerror(); return false; }
#define PGT_SYM(field, name)                                                      \
    *(void**)(&g_nccl.field) = dlsym(g_nccl.handle, name);                        \
    if (!g_nccl.field) { err = std::string("libnccl lacks ") + name; return false; }
    PGT_SYM(GetUniqueId, "ncclGetUniqueId")
    PGT_SYM(CommInitRank, "ncclCommInitRank")
    PGT_SYM(CommDestroy, "ncclCommDestroy")
    PGT_SYM(AllReduce, "ncclAllReduce")
    PGT_SYM(Broadcast, "ncclBroadcast")
    PGT_SYM(AllGather, "ncclAllGather")
    PGT_SYM(GroupStart, "ncclGroupStart")
    PGT_SYM(GroupEnd, "ncclGroupEnd")
    PGT_SYM(GetErrorString, "ncclGetErrorString")
#undef PGT_SYM
    g_nccl.ok = true;
    return true;
}

// Gauss-Legendre nodes on [-1,1] by Newton iteration in long double (independent of the oracle's tables).
void gauss_legendre_half(int n, std::vector<double>& x, std::vector<double>& w)
{
    const int m = (n + 1) / 2;
    x.assign(m, 0.0);
    w.assign(m, 0.0);
    const long double pi = 3.14159265358979323846264338327950288L;
    for (int i = 1; i <= m; ++i) {
        long double z = cosl(pi * (i - 0.25L) / (n + 0.5L));
        long double pp = 1;
        for (int it = 0; it < 100; ++it) {
            long double p1 = 1, p2 = 0;
            for (int j = 1; j <= n; ++j) {
                long double p3 = p2;
                p2 = p1;
                p1 = ((2 * j - 1) * z * p2 - (j - 1) * p3) / j;
            }
            pp = n * (z * p1 - p2) / (z * z - 1);
            long double z1 = z;
            z = z1 - p1 / pp;
            if (fabsl(z - z1) < 1e-19L) break;
        }
        if ((n & 1) && i == m) z = 0;
        {   // derivative at the converged root
            long double p1 = 1, p2 = 0;
            for (int j = 1; j <= n; ++j) {
                long double p3 = p2;
                p2 = p1;
                p1 = ((2 * j - 1) * z * p2 - (j - 1) * p3) / j;
            }
            pp = (z == 0) ? (long double)n * (-p2) / (-1) : n * (z * p1 - p2) / (z * z - 1);
        }
        // ascending non-negative abscissae, like the GSL table the reference integrates with
        x[m - i] = (double)z;
        w[m - i] = (double)(2 / ((1 - z * z) * pp * pp));
    }
}

}  // namespace

struct pgt_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaStream_t dl_stream = nullptr;      // read-backs issued by pgt_ik_download_async (overlap the next upload)
    cudaEvent_t ev_field_ready = nullptr;  // main stream: the field of the last gather is complete
    cudaEvent_t ev_dl_copy[4] = { nullptr, nullptr, nullptr, nullptr };   // dl_stream: the last four asynchronous read-backs have landed in host memory
    long long dl_seq = 0, dl_waited = 0;   // read-backs enqueued / known to have landed
    cudaEvent_t ev_dl_done = nullptr;      // dl_stream: the last asynchronous read-back no longer reads the field buffers
    bool dl_pending = false;
    std::string err;
    PgtLaunchCounter launches{ 0 };

    // models
    bool have_pot = false, have_cg = false;
    PgtPot pot{};
    PgtCG cg{};
    int nq = 0;
    double gl5[7] = { 0, 0, 0, 0, 0, 0, 0 };
    double gl_s[5] = { 0, 0, 0, 0, 0 }, gl_w[5] = { 0, 0, 0, 0, 0 };   // symmetric 5-point rule on [0,1]: h1, h1^2, h2, h2^2, w0, w1, w2
    double max_rcut = 0.0;
    double cell_side_user = 0.0;

    // frame
    bool have_frame = false, cells_built = false, bonds_built = false;
    long long N = 0;
    int is_f32 = 0;
    int dim = 2;
    PgtBox box{};
    double vol = 0.0;
    const void *src_pos = nullptr, *src_diam = nullptr, *src_mass = nullptr, *src_vel = nullptr;
    // positions and diameters are double-buffered and uploaded on their own stream: frame s+1 crosses PCIe while frame s
    // is computed on the main stream -- one context pipelines a trajectory (kernels of different frames never co-run)
    DevBuf raw_pos2[2], raw_diam2[2], raw_mass, raw_vel;
    int raw_slot = 0;                 // buffer pair of the frame last uploaded
    int raw_cur = -1;                 // buffer pair the current frame's stage 1 reads (-1: caller's device pointers)
    int ul_wait_slot = -1;            // the main stream has yet to wait for this upload
    cudaStream_t ul_stream = nullptr;
    cudaEvent_t ev_ul_done[2] = { nullptr, nullptr }, ev_raw_free[2] = { nullptr, nullptr };
    bool raw_free_valid[2] = { false, false };
    PgtCells cells{};
    int n_live = 0;
    long long n_hb = 0;
    DevBuf cell_of, cell_count, cell_fill, sorted_orig, scan_scratch, eflags;
    DevBuf sx, sy, sxy, sd, skxx, skxy, skyy;
    DevBuf sz, sm3, svx3, svy3, svz3;   // Dim = 3: wrapped z, mass and velocity components of the sorted particles
    Pgt3Cells cells3{};
    DevBuf hb_cnt, hb_b, hb_cf, hb_j, hb_pm, nb_list;
    bool hb_pm_valid = false;         // the primary masks cover every particle's records (single-search path, <= 16 bonds each)
    DevBuf gl_nodes, gl_weights;      // this context's quadrature rule (never shared between contexts)
    bool bond_single_search = true;   // option "bond_single_search"
    bool bond_fill_flat = true;       // option "bond_fill_flat": warp-flattened record writer (else thread per particle)
    double hb_cscale = 1.0;   // scale folded into the records' |b|^2 entry (1/R^2 of the CG function, or 1)

    // grid
    int grid_mode = -1;
    int nx = 0, ny = 0;
    long long G = 0, g0 = 0, g1 = 0;
    DevBuf gxs, gys, gpts;

    // outputs
    DevBuf out_virial, out_kinetic, out_density, out_elastic;
    DevBuf pack;                 // packed symmetric field for pgt_ik_download_sym_async
    DevBuf disp;                 // N x 3 displacements, original particle order
    bool have_disp = false;
    unsigned last_ik_mask = 0;
    DevBuf thermo_partials, thermo_out;
    double* h_small = nullptr;   // mapped page-locked scratch for small read-backs (written by k_publish)
    void* d_small = nullptr;     // its device alias

    // timing
    cudaEvent_t ev0[PGT_STAGE_COUNT]{}, ev1[PGT_STAGE_COUNT]{};
    bool ev_valid[PGT_STAGE_COUNT]{};

    cudaEvent_t user_ev[16]{};

    // tuning
    int ik_variant = 0;   // 0: automatic; 2: two-phase gather; 7: quad gather (where applicable)
    int quad_mode = PGT_QUAD_FIXEDPOINT;     // PGT_QUAD_ADAPTIVE: QAG / GK15 bond quadrature
    double qag_epsabs = 0.0, qag_epsrel = 0.0;
    int qag_limit = 0;
    bool grid_monotone = false;
    std::vector<double> h_xs;         // host copy of the Cartesian x axis (slab of a grid shard)
    bool slab_cull = true;            // option "slab_cull"
    bool culled = false;              // the frame in HBM holds only the particles the grid shard can reach
    double slab_lo = 0.0, slab_hi = 0.0;
    int64_t part_chunk = 0;           // pgt_frame_upload_part: particles per rank part (0: whole-frame upload)
    int thermo_part = 0, thermo_nparts = 1;
    // deferred counts (option "async_counts"): once a frame has been counted exactly, the following frames of the same
    // shape size their arrays from its counts (with a margin) and do not wait for their own; the counts and any error
    // reach the host at its next synchronisation point
    bool async_counts = true;
    bool counts_pending = false;      // n_live / n_hb below are bounds, the exact values sit in mapped memory
    bool cells_deferred = false;      // this frame's cell stage did not wait
    bool eflags_init = false;
    int known_nlive = -1;             // exact counts of the last frame counted (or resolved)
    long long known_nhb = -1;
    long long known_sig[8] = { 0, 0, 0, 0, 0, 0, 0, 0 };   // the shape those counts belong to
    long long hb_cap = 0;             // records the half-bond arrays hold
    long long field_g0 = 0, field_g1 = 0;   // the shard the fields in HBM were computed for (pgt_ik_compute)
    double span_x[4] = { -1, -1, -1, -1 }, span_y[4] = { -1, -1, -1, -1 };   // extents of the quad gather's grid tiles (cartesian grids)

    // nccl
    pgtNcclComm comm = nullptr;
    int nranks = 1, rank = 0;
    DevBuf comm_scratch;
};

namespace {

int fail(pgt_ctx* c, int code, const char* fmt, ...)
{
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    if (c) c->err = buf;
    else g_create_error = buf;
    return code;
}

#define CK(call)                                                                                       \
    do {                                                                                               \
        cudaError_t e_ = (call);                                                                       \
        if (e_ != cudaSuccess) return fail(ctx, PGT_ERR_CUDA, "%s failed: %s", #call, cudaGetErrorString(e_)); \
    } while (0)

int ensure(pgt_ctx* ctx, DevBuf& b, size_t bytes)
{
    if (bytes <= b.cap) return PGT_OK;
    if (b.p) CK(cudaFree(b.p));
    b.p = nullptr;
    b.cap = 0;
    size_t want = bytes + bytes / 8 + 256;
    CK(cudaMalloc(&b.p, want));
    b.cap = want;
    return PGT_OK;
}
#define ENSURE(buf, bytes)                          \
    do {                                            \
        int rc_ = ensure(ctx, buf, (bytes));        \
        if (rc_ != PGT_OK) return rc_;              \
    } while (0)

void stage_begin(pgt_ctx* ctx, int s) { cudaEventRecord(ctx->ev0[s], ctx->stream); }
void stage_end(pgt_ctx* ctx, int s)
{
    cudaEventRecord(ctx->ev1[s], ctx->stream);
    ctx->ev_valid[s] = true;
}

int check_eflags(pgt_ctx* ctx, int flags)
{
    if (flags & PGT_EFLAG_WRAP_RUNAWAY)
        return fail(ctx, PGT_ERR_DATA, "a particle lies more than %d box lengths outside the box", PGT_WRAP_MAX_ITERS);
    if (flags & PGT_EFLAG_Z_NONZERO)
        return fail(ctx, PGT_ERR_UNSUPPORTED, "2-D frame with non-zero z coordinates is not supported on device");
    return PGT_OK;
}

// mapped slots of the stage counts (ints): [0] n_live, [1] cell-stage flags, [2] n_hb, [3] bond-stage flags, [4] sticky
int* h_counts(const pgt_ctx* c) { return reinterpret_cast<int*>(c->h_small) + 512; }
int* d_counts(const pgt_ctx* c) { return reinterpret_cast<int*>(c->d_small) + 512; }

void frame_signature(const pgt_ctx* c, long long sig[8])
{
    sig[0] = c->N;
    memcpy(&sig[1], &c->box.L[0], 8);
    memcpy(&sig[2], &c->box.L[1], 8);
    sig[3] = ((long long)c->cells.ncx << 32) | (unsigned)c->cells.ncy;
    sig[4] = c->culled ? 1 : 0;
    memcpy(&sig[5], &c->slab_lo, 8);
    memcpy(&sig[6], &c->slab_hi, 8);
    sig[7] = c->is_f32;
}

int sticky_error(pgt_ctx* ctx, int sticky)
{
    // the context falls back to exact counting until a frame has been counted again
    ctx->known_nlive = -1;
    ctx->known_nhb = -1;
    cudaMemsetAsync(ctx->eflags.as<int>() + 2, 0, sizeof(int), ctx->stream);
    h_counts(ctx)[4] = 0;
    int rc = check_eflags(ctx, sticky);
    if (rc != PGT_OK) return rc;
    if (sticky & PGT_EFLAG_BOND_OVERFLOW)
        return fail(ctx, PGT_ERR_DATA, "a particle has more than %d bonds in a frame whose counts were deferred: submit the frames since the last synchronisation again (the context counts exactly now)", PGT_NB_SLOTS);
    return fail(ctx, PGT_ERR_DATA, "a frame's particle or bond count outgrew the previous frame's by more than the margin while its counts were deferred: submit the frames since the last synchronisation again (the context counts exactly now)");
}

// Exact counts on the host: waits for the stream when the last frame deferred them.
int resolve_counts(pgt_ctx* ctx)
{
    if (!ctx->counts_pending) return PGT_OK;
    CK(cudaStreamSynchronize(ctx->stream));
    ctx->counts_pending = false;
    const int* h = h_counts(ctx);
    if (h[4]) return sticky_error(ctx, h[4]);
    ctx->n_live = h[0];
    ctx->known_nlive = h[0];
    if (ctx->bonds_built) { ctx->n_hb = h[2]; ctx->known_nhb = h[2]; }
    return PGT_OK;
}
#define RESOLVE()                                   \
    do {                                            \
        int rc_ = resolve_counts(ctx);              \
        if (rc_ != PGT_OK) return rc_;              \
    } while (0)

PgtFrameView frame_view(const pgt_ctx* c)
{
    PgtFrameView v{};
    v.n = c->n_live;
    v.n_dev = c->counts_pending ? c->cell_count.as<int>() + (size_t)c->cells.ncx * c->cells.ncy : nullptr;
    v.sx = c->sx.as<double>();
    v.sy = c->sy.as<double>();
    v.sxy = c->sxy.as<double2>();
    v.sd = c->sd.as<double>();
    const bool kin = c->src_vel != nullptr;
    v.skxx = kin ? c->skxx.as<double>() : nullptr;
    v.skxy = kin ? c->skxy.as<double>() : nullptr;
    v.skyy = kin ? c->skyy.as<double>() : nullptr;
    v.sorig = c->sorted_orig.as<int>();
    v.cell_start = c->cell_count.as<int>();
    v.hb_off = c->hb_cnt.as<int>();
    v.hb_b = c->hb_b.as<double2>();
    v.hb_cf = c->hb_cf.as<double2>();
    v.hb_j = c->hb_j.as<int>();
    v.hb_pm = c->hb_pm_valid ? c->hb_pm.as<unsigned short>() : nullptr;
    v.n_hb = c->n_hb;
    v.hb_cscale = c->hb_cscale;
    return v;
}

int set_box(pgt_ctx* ctx, const double box[3], int dim)
{
    if (dim != 2 && dim != 3) return fail(ctx, PGT_ERR_ARG, "dim must be 2 or 3 (got %d)", dim);
    ctx->dim = dim;
    double L[3] = { box[0], box[1], dim == 2 ? 1.0 : box[2] };   // utils.py:15-16 forces Lz = 1 in 2-D
    for (int d = 0; d < 3; ++d) {
        if (!(L[d] > 0) || !std::isfinite(L[d])) return fail(ctx, PGT_ERR_ARG, "box length %d is not positive", d);
        // SimBox::setBounds with origin 0 (SimBox.cc:48-52)
        ctx->box.hi[d] = 0.5 * L[d] - 0.0;
        ctx->box.lo[d] = (-1.0) * ctx->box.hi[d] - 2.0 * 0.0;
        ctx->box.L[d] = ctx->box.hi[d] - ctx->box.lo[d];   // Aboria: bmax - bmin
    }
    ctx->box.dim = dim;
    ctx->box.pstride = 3;
    ctx->vol = dim == 2 ? L[0] * L[1] : L[0] * L[1] * L[2];    // SimBox.cc:36-44
    return PGT_OK;
}

}  // namespace

extern "C" {

int pgt_device_count(int* n)
{
    cudaError_t e = cudaGetDeviceCount(n);
    if (e != cudaSuccess) { *n = 0; g_create_error = cudaGetErrorString(e); return PGT_ERR_CUDA; }
    return PGT_OK;
}

int pgt_ctx_create(int device, pgt_ctx** out)
{
    if (!out) return PGT_ERR_ARG;
    *out = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return fail(nullptr, PGT_ERR_CUDA, "no CUDA device available (%s); libpgt_cuda has no CPU fallback",
                    e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
    if (device < 0 || device >= ndev) return fail(nullptr, PGT_ERR_ARG, "device %d out of range [0,%d)", device, ndev);
    pgt_ctx* ctx = new pgt_ctx();
    ctx->device = device;
    if ((e = cudaSetDevice(device)) != cudaSuccess || (e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking)) != cudaSuccess) {
        fail(nullptr, PGT_ERR_CUDA, "cannot initialise device %d: %s", device, cudaGetErrorString(e));
        delete ctx;
        return PGT_ERR_CUDA;
    }
    cudaStreamCreateWithFlags(&ctx->dl_stream, cudaStreamNonBlocking);
    cudaStreamCreateWithFlags(&ctx->ul_stream, cudaStreamNonBlocking);
    for (int k = 0; k < 2; ++k) {
        cudaEventCreateWithFlags(&ctx->ev_ul_done[k], cudaEventDisableTiming);
        cudaEventCreateWithFlags(&ctx->ev_raw_free[k], cudaEventDisableTiming);
    }
    cudaEventCreateWithFlags(&ctx->ev_field_ready, cudaEventDisableTiming);
    cudaEventCreateWithFlags(&ctx->ev_dl_done, cudaEventDisableTiming);
    for (auto& e : ctx->ev_dl_copy) cudaEventCreateWithFlags(&e, cudaEventDisableTiming);
    for (int s = 0; s < PGT_STAGE_COUNT; ++s) {
        cudaEventCreate(&ctx->ev0[s]);
        cudaEventCreate(&ctx->ev1[s]);
    }
    cudaHostAlloc((void**)&ctx->h_small, 4096, cudaHostAllocMapped);
    cudaHostGetDevicePointer(&ctx->d_small, ctx->h_small, 0);
    if (ctx->h_small) memset(ctx->h_small, 0, 4096);
    if (const char* env = getenv("PGT_ASYNC_COUNTS")) ctx->async_counts = atoi(env) != 0;
    if (const char* env = getenv("PGT_BOND_FILL_FLAT")) ctx->bond_fill_flat = atoi(env) != 0;
    if (const char* env = getenv("PGT_IK_VARIANT")) {
        int l = atoi(env);
        if (l == 0 || l == 2 || l == 7) ctx->ik_variant = l;
    }
    *out = ctx;
    return PGT_OK;
}

int pgt_ctx_destroy(pgt_ctx* ctx)
{
    if (!ctx) return PGT_OK;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    if (ctx->ul_stream) cudaStreamSynchronize(ctx->ul_stream);
    if (ctx->dl_stream) cudaStreamSynchronize(ctx->dl_stream);
    if (ctx->comm && g_nccl.ok) g_nccl.CommDestroy(ctx->comm);
    DevBuf* bufs[] = { &ctx->raw_pos2[0], &ctx->raw_pos2[1], &ctx->raw_diam2[0], &ctx->raw_diam2[1], &ctx->raw_mass, &ctx->raw_vel, &ctx->cell_of, &ctx->cell_count, &ctx->cell_fill,
                       &ctx->sorted_orig, &ctx->scan_scratch, &ctx->eflags, &ctx->sx, &ctx->sy, &ctx->sxy, &ctx->sd, &ctx->skxx, &ctx->skxy,
                       &ctx->skyy, &ctx->sz, &ctx->sm3, &ctx->svx3, &ctx->svy3, &ctx->svz3, &ctx->hb_cnt, &ctx->hb_b, &ctx->hb_cf, &ctx->hb_j, &ctx->hb_pm, &ctx->nb_list, &ctx->gl_nodes,
                       &ctx->gl_weights, &ctx->gxs, &ctx->gys, &ctx->gpts,
                       &ctx->out_virial, &ctx->out_kinetic, &ctx->out_density, &ctx->out_elastic, &ctx->disp, &ctx->pack, &ctx->thermo_partials,
                       &ctx->thermo_out,
                       &ctx->comm_scratch };
    for (DevBuf* b : bufs)
        if (b->p) cudaFree(b->p);
    for (int s = 0; s < PGT_STAGE_COUNT; ++s) {
        cudaEventDestroy(ctx->ev0[s]);
        cudaEventDestroy(ctx->ev1[s]);
    }
    if (ctx->h_small) cudaFreeHost(ctx->h_small);
    for (auto& e : ctx->user_ev)
        if (e) cudaEventDestroy(e);
    if (ctx->ev_field_ready) cudaEventDestroy(ctx->ev_field_ready);
    if (ctx->ev_dl_done) cudaEventDestroy(ctx->ev_dl_done);
    for (auto& e : ctx->ev_dl_copy)
        if (e) cudaEventDestroy(e);
    if (ctx->dl_stream) cudaStreamDestroy(ctx->dl_stream);
    if (ctx->ul_stream) cudaStreamDestroy(ctx->ul_stream);
    for (int k = 0; k < 2; ++k) {
        if (ctx->ev_ul_done[k]) cudaEventDestroy(ctx->ev_ul_done[k]);
        if (ctx->ev_raw_free[k]) cudaEventDestroy(ctx->ev_raw_free[k]);
    }
    cudaStreamDestroy(ctx->stream);
    delete ctx;
    return PGT_OK;
}

const char* pgt_last_error(const pgt_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

int pgt_synchronize(pgt_ctx* ctx)
{
    if (!ctx) return PGT_ERR_ARG;
    CK(cudaStreamSynchronize(ctx->stream));
    if (ctx->dl_pending) {
        CK(cudaStreamSynchronize(ctx->dl_stream));
        ctx->dl_pending = false;
    }
    return resolve_counts(ctx);                              // deferred counts and errors of the frames in flight
}

void* pgt_stream(pgt_ctx* ctx) { return ctx ? (void*)ctx->stream : nullptr; }

int64_t pgt_kernel_launches(const pgt_ctx* ctx) { return ctx ? ctx->launches.n : 0; }

// ---------------------------------------------------------------------------------- models
static int expected_nparams(int model)
{
    switch (model) {
    case PGT_POT_LJ: case PGT_POT_FORCESHIFTED_LJ: return 2;
    case PGT_POT_POLY10: case PGT_POT_POLYYUKAWA: return 6;
    case PGT_POT_POLY12: case PGT_POT_POLY18: case PGT_POT_POLY106: case PGT_POT_POLYLJ: return 5;
    }
    return -1;
}

static int fill_pot(PgtPot& P, int model_id, double scaled_rcut, const double* params, int nparams)
{
    const int want = expected_nparams(model_id);
    if (want < 0 || nparams < want || nparams > 8 || !params) return -1;
    P.model = model_id;
    P.nparams = nparams;
    P.scaled_rcut = scaled_rcut;
    P.rcutsq = scaled_rcut * scaled_rcut;
    for (int k = 0; k < 8; ++k) P.p[k] = k < nparams ? params[k] : 0.0;
    return 0;
}

int pgt_potential_set(pgt_ctx* ctx, int model_id, double scaled_rcut, const double* params, int nparams)
{
    if (!ctx) return PGT_ERR_ARG;
    if (fill_pot(ctx->pot, model_id, scaled_rcut, params, nparams) != 0)
        return fail(ctx, PGT_ERR_ARG, "bad pair model %d / parameter count %d", model_id, nparams);
    ctx->have_pot = true;
    ctx->bonds_built = false;
    return PGT_OK;
}

int pgt_cg_set(pgt_ctx* ctx, int kernel_id, double cg_rcut, int quad_mode, int order)
{
    if (!ctx) return PGT_ERR_ARG;
    if (kernel_id < PGT_CG_OCTIC || kernel_id > PGT_CG_MOLLIFIER) return fail(ctx, PGT_ERR_ARG, "unknown CG kernel %d", kernel_id);
    if (quad_mode != PGT_QUAD_FIXEDPOINT)
        return fail(ctx, PGT_ERR_ARG, "adaptive bond quadrature takes tolerances: use pgt_cg_set_adaptive");
    if (order < 1 || order > PGT_MAX_GL) return fail(ctx, PGT_ERR_UNSUPPORTED, "quadrature order %d outside [1,%d]", order, PGT_MAX_GL);
    if (!(cg_rcut > 0)) return fail(ctx, PGT_ERR_ARG, "cg_rcut must be positive");
    CK(cudaSetDevice(ctx->device));
    ctx->cg.kernel = kernel_id;
    ctx->cg.order = order;
    ctx->cg.rcut = cg_rcut;
    ctx->cg.rcutsq = cg_rcut * cg_rcut;
    ctx->cg.norm = pgt_cg_norm(kernel_id, cg_rcut);
    // nodes in GSL's visiting order: centre (odd n), then +-pairs; weights carry (b-a)/2 = 0.5
    std::vector<double> x, w, s, ww;
    gauss_legendre_half(order, x, w);
    const int m = (order + 1) / 2;
    int i0 = 0;
    if (order & 1) { s.push_back(0.5); ww.push_back(0.5 * w[0]); i0 = 1; }
    for (int i = i0; i < m; ++i) {
        const double Ax = 0.5 * x[i];
        s.push_back(0.5 + Ax); ww.push_back(0.5 * w[i]);
        s.push_back(0.5 - Ax); ww.push_back(0.5 * w[i]);
    }
    ctx->nq = (int)s.size();
    if (order == 5) {
        const double h1 = 0.5 * x[1], h2 = 0.5 * x[2];
        const double g[7] = { h1, h1 * h1, h2, h2 * h2, 0.5 * w[0], 0.5 * w[1], 0.5 * w[2] };
        for (int k = 0; k < 7; ++k) ctx->gl5[k] = g[k];
    }
    if (ctx->nq == 5)
        for (int k = 0; k < 5; ++k) { ctx->gl_s[k] = s[k]; ctx->gl_w[k] = ww[k]; }
    ENSURE(ctx->gl_nodes, sizeof(double) * PGT_MAX_GL);
    ENSURE(ctx->gl_weights, sizeof(double) * PGT_MAX_GL);
    CK(cudaMemcpyAsync(ctx->gl_nodes.p, s.data(), sizeof(double) * ctx->nq, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(ctx->gl_weights.p, ww.data(), sizeof(double) * ctx->nq, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));   // s/ww are stack-owned
    ctx->quad_mode = PGT_QUAD_FIXEDPOINT;
    ctx->have_cg = true;
    return PGT_OK;
}

int pgt_cg_set_adaptive(pgt_ctx* ctx, int kernel_id, double cg_rcut, double epsabs, double epsrel, int limit)
{
    if (!ctx) return PGT_ERR_ARG;
    if (kernel_id < PGT_CG_OCTIC || kernel_id > PGT_CG_MOLLIFIER) return fail(ctx, PGT_ERR_ARG, "unknown CG kernel %d", kernel_id);
    if (!(cg_rcut > 0)) return fail(ctx, PGT_ERR_ARG, "cg_rcut must be positive");
    // gsl_integration_qag: "tolerance cannot be achieved with given epsabs and epsrel" (GSL_EBADTOL), limit > workspace
    if (epsabs <= 0 && (epsrel < 50 * 2.2204460492503131e-16 || epsrel < 0.5e-28))
        return fail(ctx, PGT_ERR_ARG, "tolerance cannot be achieved with given epsabs and epsrel");
    if (limit < 1) return fail(ctx, PGT_ERR_ARG, "maxiter must be >= 1");
    // the fixed-point tables keep a valid (1-point) rule so that stages that only need Delta() are unaffected
    const int rc = pgt_cg_set(ctx, kernel_id, cg_rcut, PGT_QUAD_FIXEDPOINT, 1);
    if (rc != PGT_OK) return rc;
    ctx->quad_mode = PGT_QUAD_ADAPTIVE;
    ctx->qag_epsabs = epsabs;
    ctx->qag_epsrel = epsrel;
    ctx->qag_limit = limit;
    return PGT_OK;
}

int pgt_set_max_rcut(pgt_ctx* ctx, double max_rcut)
{
    if (!ctx) return PGT_ERR_ARG;
    if (!(max_rcut > 0) || !std::isfinite(max_rcut)) return fail(ctx, PGT_ERR_ARG, "max_rcut must be positive");
    if (max_rcut != ctx->max_rcut) { ctx->cells_built = false; ctx->bonds_built = false; }
    ctx->max_rcut = max_rcut;
    return PGT_OK;
}

int pgt_set_cell_size(pgt_ctx* ctx, double cell_side)
{
    if (!ctx) return PGT_ERR_ARG;
    ctx->cell_side_user = cell_side > 0 ? cell_side : 0.0;
    ctx->cells_built = false;
    ctx->bonds_built = false;
    return PGT_OK;
}

// ---------------------------------------------------------------------------------- frame
static int set_frame_common(pgt_ctx* ctx, int is_f32, int64_t N, const double box[3], int dim)
{
    if (N < 0 || N > 2000000000LL) return fail(ctx, PGT_ERR_ARG, "particle count %lld out of range", (long long)N);
    int rc = set_box(ctx, box, dim);
    if (rc != PGT_OK) return rc;
    ctx->N = N;
    ctx->is_f32 = is_f32 ? 1 : 0;
    ctx->have_frame = true;
    ctx->cells_built = false;
    ctx->bonds_built = false;
    return PGT_OK;
}

// The buffer pair of the next upload: the other one than the frame in flight uses.  The upload stream waits until the
// last reader of that pair (stage 1 of the frame before last) is done.
static int upload_slot(pgt_ctx* ctx, size_t pos_bytes, size_t diam_bytes)
{
    const int k = ctx->raw_slot ^ 1;
    if (pos_bytes > ctx->raw_pos2[k].cap || diam_bytes > ctx->raw_diam2[k].cap) cudaStreamSynchronize(ctx->stream);   // re-allocation
    int rc = ensure(ctx, ctx->raw_pos2[k], pos_bytes);
    if (rc == PGT_OK) rc = ensure(ctx, ctx->raw_diam2[k], diam_bytes);
    if (rc != PGT_OK) return -1;
    if (ctx->raw_free_valid[k]) cudaStreamWaitEvent(ctx->ul_stream, ctx->ev_raw_free[k], 0);
    ctx->raw_slot = k;
    return k;
}

static void upload_done(pgt_ctx* ctx, int k)
{
    cudaEventRecord(ctx->ev_ul_done[k], ctx->ul_stream);
    ctx->ul_wait_slot = k;
    ctx->raw_cur = k;
    ctx->src_pos = ctx->raw_pos2[k].p;
    ctx->src_diam = ctx->raw_diam2[k].p;
}

// The main stream joins the upload of the frame it is about to use.
static void wait_upload(pgt_ctx* ctx)
{
    if (ctx->ul_wait_slot >= 0) {
        cudaStreamWaitEvent(ctx->stream, ctx->ev_ul_done[ctx->ul_wait_slot], 0);
        ctx->ul_wait_slot = -1;
    }
}

static int frame_upload_impl(pgt_ctx* ctx, const void* pos, const void* diam, const void* mass, const void* vel, int is_f32,
                             int64_t N, const double box[3], int dim, int pstride)
{
    if (!ctx || !pos || !diam || !box) return fail(ctx, PGT_ERR_ARG, "pgt_frame_upload: null argument");
    CK(cudaSetDevice(ctx->device));
    int rc = set_frame_common(ctx, is_f32, N, box, dim);
    if (rc != PGT_OK) return rc;
    ctx->box.pstride = pstride;
    const size_t es = is_f32 ? 4 : 8;
    const int k = upload_slot(ctx, (size_t)N * pstride * es, (size_t)N * es);
    if (k < 0) return PGT_ERR_CUDA;
    cudaEventRecord(ctx->ev0[PGT_STAGE_UPLOAD], ctx->ul_stream);
    CK(cudaMemcpyAsync(ctx->raw_pos2[k].p, pos, (size_t)N * pstride * es, cudaMemcpyHostToDevice, ctx->ul_stream));
    CK(cudaMemcpyAsync(ctx->raw_diam2[k].p, diam, (size_t)N * es, cudaMemcpyHostToDevice, ctx->ul_stream));
    cudaEventRecord(ctx->ev1[PGT_STAGE_UPLOAD], ctx->ul_stream);
    ctx->ev_valid[PGT_STAGE_UPLOAD] = true;
    upload_done(ctx, k);
    ctx->src_mass = nullptr;
    ctx->src_vel = nullptr;
    ctx->have_disp = false;
    if (mass) {
        ENSURE(ctx->raw_mass, (size_t)N * es);
        CK(cudaMemcpyAsync(ctx->raw_mass.p, mass, (size_t)N * es, cudaMemcpyHostToDevice, ctx->stream));
        ctx->src_mass = ctx->raw_mass.p;
    }
    if (vel) {
        ENSURE(ctx->raw_vel, (size_t)N * 3 * es);
        CK(cudaMemcpyAsync(ctx->raw_vel.p, vel, (size_t)N * 3 * es, cudaMemcpyHostToDevice, ctx->stream));
        ctx->src_vel = ctx->raw_vel.p;
    }
    ctx->part_chunk = 0;
    return PGT_OK;
}

int pgt_frame_upload(pgt_ctx* ctx, const void* pos, const void* diam, const void* mass, const void* vel, int is_f32,
                     int64_t N, const double box[3], int dim)
{
    return frame_upload_impl(ctx, pos, diam, mass, vel, is_f32, N, box, dim, 3);
}

int pgt_frame_upload_xy(pgt_ctx* ctx, const void* pos_xy, const void* diam, const void* mass, const void* vel, int is_f32,
                        int64_t N, const double box[3])
{
    return frame_upload_impl(ctx, pos_xy, diam, mass, vel, is_f32, N, box, 2, 2);
}

int pgt_frame_upload_part(pgt_ctx* ctx, const void* pos_part, const void* diam_part, int is_f32, int64_t first, int64_t count,
                          int64_t N, const double box[3], int dim)
{
    if (!ctx || !box || (count > 0 && (!pos_part || !diam_part))) return fail(ctx, PGT_ERR_ARG, "pgt_frame_upload_part: null argument");
    if (first < 0 || count < 0 || first + count > N) return fail(ctx, PGT_ERR_ARG, "pgt_frame_upload_part: part [%lld, %lld) outside [0, %lld)", (long long)first, (long long)(first + count), (long long)N);
    const int nr = std::max(ctx->nranks, 1);
    const int64_t chunk = (N + nr - 1) / nr;
    if (ctx->comm && (first != (int64_t)ctx->rank * chunk || count != std::max<int64_t>(0, std::min(chunk, N - first))))
        return fail(ctx, PGT_ERR_ARG, "pgt_frame_upload_part: rank %d of %d owns particles [%lld, %lld)", ctx->rank, nr,
                    (long long)(ctx->rank * chunk), (long long)std::min<int64_t>(N, (ctx->rank + 1) * chunk));
    CK(cudaSetDevice(ctx->device));
    int rc = set_frame_common(ctx, is_f32, N, box, dim);
    if (rc != PGT_OK) return rc;
    const size_t es = is_f32 ? 4 : 8;
    // room for nranks equal parts: ncclAllGather moves equal counts
    const int k = upload_slot(ctx, (size_t)chunk * nr * 3 * es, (size_t)chunk * nr * es);
    if (k < 0) return PGT_ERR_CUDA;
    cudaEventRecord(ctx->ev0[PGT_STAGE_UPLOAD], ctx->ul_stream);
    if (count > 0) {
        CK(cudaMemcpyAsync(static_cast<char*>(ctx->raw_pos2[k].p) + (size_t)first * 3 * es, pos_part, (size_t)count * 3 * es, cudaMemcpyHostToDevice, ctx->ul_stream));
        CK(cudaMemcpyAsync(static_cast<char*>(ctx->raw_diam2[k].p) + (size_t)first * es, diam_part, (size_t)count * es, cudaMemcpyHostToDevice, ctx->ul_stream));
    }
    cudaEventRecord(ctx->ev1[PGT_STAGE_UPLOAD], ctx->ul_stream);
    ctx->ev_valid[PGT_STAGE_UPLOAD] = true;
    upload_done(ctx, k);
    ctx->src_mass = nullptr;
    ctx->src_vel = nullptr;
    ctx->have_disp = false;
    ctx->part_chunk = chunk;
    return PGT_OK;
}

int pgt_frame_set_device(pgt_ctx* ctx, const void* d_pos, const void* d_diam, const void* d_mass, const void* d_vel, int is_f32,
                         int64_t N, const double box[3], int dim)
{
    if (!ctx || !d_pos || !d_diam || !box) return fail(ctx, PGT_ERR_ARG, "pgt_frame_set_device: null argument");
    int rc = set_frame_common(ctx, is_f32, N, box, dim);
    if (rc != PGT_OK) return rc;
    ctx->src_pos = d_pos;
    ctx->src_diam = d_diam;
    ctx->raw_cur = -1;
    ctx->ul_wait_slot = -1;
    ctx->src_mass = d_mass;
    ctx->src_vel = d_vel;
    ctx->have_disp = false;
    return PGT_OK;
}

// The x interval of the particles a grid shard can reach (Cartesian grid, non-decreasing axes): the shard's grid columns,
// widened to whole gather tiles (32 columns: a warp's control flow, and so a point's summation order, depends on its
// tile-mates), plus the coarse-graining radius, the cell rounding of a tile's rectangle and the reach of the bond search.
static bool shard_slab(const pgt_ctx* ctx, const PgtCells& c, double* lo, double* hi)
{
    if (!ctx->slab_cull || ctx->grid_mode != 0 || !ctx->grid_monotone || !ctx->have_cg || ctx->dim != 2) return false;
    if (ctx->nx <= 0 || ctx->ny <= 0 || ctx->g1 <= ctx->g0 || (ctx->g0 == 0 && ctx->g1 == ctx->G)) return false;
    if ((int)ctx->h_xs.size() != ctx->nx) return false;
    const int i_lo = (int)(ctx->g0 / ctx->ny), i_hi = (int)((ctx->g1 - 1) / ctx->ny);
    const int a_lo = (i_lo / 32) * 32, a_hi = std::min(ctx->nx - 1, (i_hi / 32) * 32 + 31);
    const double cw = 1.0 / c.inv_cwx;
    const double margin = ctx->cg.rcut + (c.reach + 4) * cw;
    *lo = ctx->h_xs[a_lo] - margin;
    *hi = ctx->h_xs[a_hi] + margin;
    return *hi - *lo < ctx->box.L[0];
}

// ---------------------------------------------------------------------------------- Dim = 3 (pgt_dim3.cu)
// The reference's 3-D typedefs (module-irvingkirkwood.cpp:8-13, module-thermo.cpp:7-20): parity kernels, exact counts, no
// deferral, no slab cull, no bond records (the gather runs the reference's nested search itself).
static Pgt3Frame frame3(const pgt_ctx* c)
{
    Pgt3Frame f{};
    f.n = c->n_live;
    f.sx = c->sx.as<double>(); f.sy = c->sy.as<double>(); f.sz = c->sz.as<double>(); f.sd = c->sd.as<double>();
    const bool kin = c->src_vel != nullptr;
    f.sm = kin ? c->sm3.as<double>() : nullptr;
    f.svx = kin ? c->svx3.as<double>() : nullptr;
    f.svy = kin ? c->svy3.as<double>() : nullptr;
    f.svz = kin ? c->svz3.as<double>() : nullptr;
    f.sorig = c->sorted_orig.as<int>();
    f.cell_start = c->cell_count.as<int>();
    return f;
}

static int build_cells3(pgt_ctx* ctx)
{
    const long long N = ctx->N;
    const double L[3] = { ctx->box.L[0], ctx->box.L[1], ctx->box.L[2] };
    double side = ctx->cell_side_user > 0 ? ctx->cell_side_user : ctx->max_rcut;
    side = std::max(side, std::cbrt(L[0] * L[1] * L[2] / (4.0 * (double)std::max<long long>(N, 1) + 64.0)));
    Pgt3Cells& c = ctx->cells3;
    for (;;) {
        c.ncx = std::max(1, (int)std::floor(L[0] / side));
        c.ncy = std::max(1, (int)std::floor(L[1] / side));
        c.ncz = std::max(1, (int)std::floor(L[2] / side));
        if ((long long)c.ncx * c.ncy * c.ncz <= (1LL << 27)) break;
        side *= 1.26;
    }
    c.inv_cw[0] = 1.0 / (L[0] / c.ncx);
    c.inv_cw[1] = 1.0 / (L[1] / c.ncy);
    c.inv_cw[2] = 1.0 / (L[2] / c.ncz);
    const int ncells = c.ncx * c.ncy * c.ncz;

    wait_upload(ctx);
    stage_begin(ctx, PGT_STAGE_CELLS);
    ENSURE(ctx->cell_of, sizeof(int) * (size_t)std::max<long long>(N, 1));
    ENSURE(ctx->cell_count, sizeof(int) * ((size_t)ncells + 1));
    ENSURE(ctx->cell_fill, sizeof(int) * (size_t)std::max<long long>(N, 1));
    ENSURE(ctx->sorted_orig, sizeof(int) * (size_t)std::max<long long>(N, 1));
    ENSURE(ctx->scan_scratch, sizeof(int) * (size_t)pgt_scan_scratch_ints(std::max<long long>(ncells, N) + 1));
    ENSURE(ctx->eflags, sizeof(int) * 4);
    CK(cudaMemsetAsync(ctx->cell_count.p, 0, sizeof(int) * ((size_t)ncells + 1), ctx->stream));
    CK(cudaMemsetAsync(ctx->eflags.p, 0, sizeof(int) * 4, ctx->stream));
    ctx->eflags_init = true;
    pgt3_launch_wrap_count(ctx->src_pos, ctx->is_f32, N, ctx->box, c, ctx->cell_of.as<int>(), ctx->cell_fill.as<int>(),
                           ctx->cell_count.as<int>(), ctx->eflags.as<int>(), ctx->stream, &ctx->launches);
    pgt_launch_scan_exclusive(ctx->cell_count.as<int>(), ctx->cell_count.as<int>(), (long long)ncells + 1, ctx->scan_scratch.as<int>(),
                              ctx->stream, &ctx->launches);
    pgt_launch_scatter_ids(ctx->cell_of.as<int>(), ctx->cell_fill.as<int>(), N, ctx->cell_count.as<int>(), ctx->sorted_orig.as<int>(),
                           ctx->stream, &ctx->launches);
    pgt_launch_publish(ctx->cell_count.as<int>() + ncells, 1, ctx->eflags.p, 1, ctx->d_small, ctx->stream, &ctx->launches);
    CK(cudaStreamSynchronize(ctx->stream));
    const int* h = reinterpret_cast<const int*>(ctx->h_small);
    int rc = check_eflags(ctx, h[1]);
    if (rc != PGT_OK) return rc;
    ctx->n_live = h[0];
    ctx->counts_pending = false;
    ctx->cells_deferred = false;
    ctx->known_nlive = -1;
    ctx->known_nhb = -1;
    ctx->culled = false;
    const size_t nl = (size_t)std::max(ctx->n_live, 1);
    ENSURE(ctx->sx, sizeof(double) * nl);
    ENSURE(ctx->sy, sizeof(double) * nl);
    ENSURE(ctx->sz, sizeof(double) * nl);
    ENSURE(ctx->sd, sizeof(double) * nl);
    const bool kin = ctx->src_vel != nullptr;
    if (kin) {
        ENSURE(ctx->sm3, sizeof(double) * nl);
        ENSURE(ctx->svx3, sizeof(double) * nl);
        ENSURE(ctx->svy3, sizeof(double) * nl);
        ENSURE(ctx->svz3, sizeof(double) * nl);
    }
    Pgt3FrameOut o{ ctx->sx.as<double>(), ctx->sy.as<double>(), ctx->sz.as<double>(), ctx->sd.as<double>(),
                    kin ? ctx->sm3.as<double>() : nullptr, kin ? ctx->svx3.as<double>() : nullptr,
                    kin ? ctx->svy3.as<double>() : nullptr, kin ? ctx->svz3.as<double>() : nullptr };
    pgt3_launch_sort_gather(ctx->src_pos, ctx->src_diam, ctx->src_mass, ctx->src_vel, ctx->is_f32, ctx->box, ctx->sorted_orig.as<int>(),
                            ctx->cell_count.as<int>(), ncells, o, ctx->stream, &ctx->launches);
    CK(cudaGetLastError());
    if (ctx->raw_cur >= 0) {
        CK(cudaEventRecord(ctx->ev_raw_free[ctx->raw_cur], ctx->stream));
        ctx->raw_free_valid[ctx->raw_cur] = true;
    }
    stage_end(ctx, PGT_STAGE_CELLS);
    ctx->cells_built = true;
    ctx->bonds_built = false;
    return PGT_OK;
}

static int ensure_disp(pgt_ctx* ctx);

static int ik_compute3(pgt_ctx* ctx, unsigned obs_mask)
{
    for (int d = 0; d < 3; ++d)
        if (ctx->cg.rcut >= ctx->box.L[d]) return fail(ctx, PGT_ERR_DATA, "cg_rcut %g is not smaller than the box", ctx->cg.rcut);
    if ((obs_mask & (PGT_OBS_VIRIAL | PGT_OBS_ELASTIC)) && !ctx->have_pot) return fail(ctx, PGT_ERR_STATE, "pgt_ik_compute: no pair potential");
    CK(cudaSetDevice(ctx->device));
    const long long gl = ctx->g1 - ctx->g0;
    ctx->field_g0 = ctx->g0;
    ctx->field_g1 = ctx->g1;
    const size_t n = (size_t)std::max<long long>(gl, 1);
    if (obs_mask & PGT_OBS_VIRIAL) ENSURE(ctx->out_virial, sizeof(double) * 9 * n);
    if (obs_mask & PGT_OBS_KINETIC) ENSURE(ctx->out_kinetic, sizeof(double) * 9 * n);
    if (obs_mask & PGT_OBS_DENSITY) ENSURE(ctx->out_density, sizeof(double) * n);
    if (obs_mask & PGT_OBS_ELASTIC) {
        ENSURE(ctx->out_elastic, sizeof(double) * 9 * n);
        const int rc = ensure_disp(ctx);
        if (rc != PGT_OK) return rc;
    }
    if (ctx->dl_pending) CK(cudaStreamWaitEvent(ctx->stream, ctx->ev_dl_done, 0));
    stage_begin(ctx, PGT_STAGE_IK);
    Pgt3IKArgs a{};
    a.fr = frame3(ctx);
    a.grid.mode = ctx->grid_mode;
    a.grid.nx = ctx->nx; a.grid.ny = ctx->ny;
    a.grid.xs = ctx->gxs.as<double>(); a.grid.ys = ctx->gys.as<double>(); a.grid.pts = ctx->gpts.as<double>();
    a.grid.g0 = ctx->g0; a.grid.g1 = ctx->g1;
    a.box = ctx->box;
    a.cells = ctx->cells3;
    a.cg = ctx->cg;
    a.nq = ctx->nq;
    a.gl_nodes = ctx->gl_nodes.as<double>();
    a.gl_weights = ctx->gl_weights.as<double>();
    a.adaptive = ctx->quad_mode == PGT_QUAD_ADAPTIVE ? 1 : 0;
    a.qag_epsabs = ctx->qag_epsabs; a.qag_epsrel = ctx->qag_epsrel; a.qag_limit = ctx->qag_limit;
    a.max_rcut = ctx->max_rcut;
    a.disp = (obs_mask & PGT_OBS_ELASTIC) ? ctx->disp.as<double>() : nullptr;
    a.out_virial = ctx->out_virial.as<double>();
    a.out_kinetic = ctx->out_kinetic.as<double>();
    a.out_density = ctx->out_density.as<double>();
    a.out_elastic = ctx->out_elastic.as<double>();
    a.obs_mask = obs_mask;
    if ((obs_mask & PGT_OBS_KINETIC) && !a.fr.sm && gl > 0) CK(cudaMemsetAsync(ctx->out_kinetic.p, 0, sizeof(double) * 9 * n, ctx->stream));
    CK(cudaMemsetAsync(ctx->eflags.p, 0, sizeof(int), ctx->stream));
    if (gl > 0) CK(pgt3_launch_ik(ctx->pot, a, ctx->eflags.as<int>(), ctx->stream, &ctx->launches));
    stage_end(ctx, PGT_STAGE_IK);
    if (a.adaptive) {
        pgt_launch_publish(ctx->eflags.p, 1, nullptr, 0, ctx->d_small, ctx->stream, &ctx->launches);
        CK(cudaStreamSynchronize(ctx->stream));
        const int fl = reinterpret_cast<int*>(ctx->h_small)[0];
        if (fl & PGT_EFLAG_QAG_CAPACITY)
            return fail(ctx, PGT_ERR_UNSUPPORTED, "adaptive bond quadrature: a bond needs more than %d intervals on device", PGT_QAG_CAP);
        if (fl & PGT_EFLAG_QAG_FAILED)
            return fail(ctx, PGT_ERR_DATA, "adaptive bond quadrature did not reach the tolerance within maxiter = %d intervals "
                        "(GSL's error handler aborts the reference here)", ctx->qag_limit);
    }
    ctx->last_ik_mask = obs_mask;
    return PGT_OK;
}

// out: virial 9 | kinetic 9 | energy | born 81 (100 doubles), or the elastic virial's 9 when elastic_only
static int thermo3(pgt_ctx* ctx, unsigned obs_mask, bool elastic_only, double* out)
{
    CK(cudaSetDevice(ctx->device));
    if (elastic_only) {
        const int rc = ensure_disp(ctx);
        if (rc != PGT_OK) return rc;
        obs_mask = PGT_OBS_ELASTIC;
    } else obs_mask &= ~PGT_OBS_ELASTIC;
    const int blocks = pgt3_thermo_blocks(ctx->n_live);
    const int p_begin = (int)((long long)ctx->n_live * ctx->thermo_part / ctx->thermo_nparts);
    const int p_end = (int)((long long)ctx->n_live * (ctx->thermo_part + 1) / ctx->thermo_nparts);
    ENSURE(ctx->thermo_partials, sizeof(double) * pgt3_thermo_nv() * (size_t)blocks);
    ENSURE(ctx->thermo_out, sizeof(double) * 128);
    stage_begin(ctx, PGT_STAGE_THERMO);
    CK(pgt3_launch_thermo(ctx->pot, frame3(ctx), ctx->box, ctx->cells3, ctx->max_rcut, obs_mask, elastic_only ? ctx->disp.as<double>() : nullptr,
                          1 / ctx->vol, ctx->thermo_partials.as<double>(), blocks, ctx->thermo_out.as<double>(), p_begin, p_end, ctx->stream,
                          &ctx->launches));
    stage_end(ctx, PGT_STAGE_THERMO);
    CK(cudaMemcpyAsync(ctx->h_small, ctx->thermo_out.p, sizeof(double) * 109, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    if (elastic_only) for (int k = 0; k < 9; ++k) out[k] = ctx->h_small[100 + k];
    else for (int k = 0; k < 100; ++k) out[k] = ctx->h_small[k];
    return PGT_OK;
}

int pgt_build_cells(pgt_ctx* ctx)
{
    if (!ctx) return PGT_ERR_ARG;
    if (!ctx->have_frame) return fail(ctx, PGT_ERR_STATE, "pgt_build_cells: no frame");
    if (!(ctx->max_rcut > 0)) return fail(ctx, PGT_ERR_STATE, "pgt_build_cells: max_rcut not set");
    CK(cudaSetDevice(ctx->device));
    if (ctx->dim == 3) return build_cells3(ctx);
    const long long N = ctx->N;

    // cell geometry: side ~ max_rcut/2, never more than ~8 cells per particle
    double side = ctx->cell_side_user > 0 ? ctx->cell_side_user : 0.5 * ctx->max_rcut;
    const double Lx = ctx->box.L[0], Ly = ctx->box.L[1];
    const double min_side = std::sqrt(Lx * Ly / (8.0 * (double)std::max<long long>(N, 1) + 64.0));
    side = std::max(side, min_side);
    PgtCells& c = ctx->cells;
    c.ncx = std::max(1, (int)std::floor(Lx / side));
    c.ncy = std::max(1, (int)std::floor(Ly / side));
    if ((long long)c.ncx * c.ncy > 1500000000LL) return fail(ctx, PGT_ERR_DATA, "cell grid %d x %d too large", c.ncx, c.ncy);
    const double cwx = Lx / c.ncx, cwy = Ly / c.ncy;
    c.inv_cwx = 1.0 / cwx;
    c.inv_cwy = 1.0 / cwy;
    c.reach = (int)std::ceil(ctx->max_rcut / std::min(cwx, cwy) + 1e-9);
    if (c.ncx < 2 * c.reach + 1 || c.ncy < 2 * c.reach + 1)
        return fail(ctx, PGT_ERR_DATA, "box %g x %g too small for max_rcut %g (needs at least %d cells per axis)", Lx, Ly,
                    ctx->max_rcut, 2 * c.reach + 1);
    const int ncells = c.ncx * c.ncy;

    wait_upload(ctx);
    stage_begin(ctx, PGT_STAGE_CELLS);
    ENSURE(ctx->cell_of, sizeof(int) * (size_t)std::max<long long>(N, 1));
    ENSURE(ctx->cell_count, sizeof(int) * ((size_t)ncells + 1));
    ENSURE(ctx->cell_fill, sizeof(int) * (size_t)std::max<long long>(N, 1));   // per particle: arrival rank in its cell
    ENSURE(ctx->sorted_orig, sizeof(int) * (size_t)std::max<long long>(N, 1));
    ENSURE(ctx->scan_scratch, sizeof(int) * (size_t)pgt_scan_scratch_ints(std::max<long long>(ncells, N) + 1));
    ENSURE(ctx->eflags, sizeof(int) * 4);
    CK(cudaMemsetAsync(ctx->cell_count.p, 0, sizeof(int) * ((size_t)ncells + 1), ctx->stream));
    if (!ctx->eflags_init) { CK(cudaMemsetAsync(ctx->eflags.p, 0, sizeof(int) * 4, ctx->stream)); ctx->eflags_init = true; }
    CK(cudaMemsetAsync(ctx->eflags.p, 0, sizeof(int) * 2, ctx->stream));      // [0] stage flags; [2] is the sticky word

    double slab_lo = 0.0, slab_hi = 0.0;
    ctx->culled = shard_slab(ctx, c, &slab_lo, &slab_hi);
    ctx->slab_lo = slab_lo;
    ctx->slab_hi = slab_hi;
    pgt_launch_wrap_count(ctx->src_pos, ctx->is_f32, N, ctx->box, c, ctx->cell_of.as<int>(), ctx->cell_fill.as<int>(), ctx->cell_count.as<int>(),
                          ctx->eflags.as<int>(), ctx->culled ? 1 : 0, slab_lo, slab_hi, ctx->stream, &ctx->launches);
    pgt_launch_scan_exclusive(ctx->cell_count.as<int>(), ctx->cell_count.as<int>(), (long long)ncells + 1,
                              ctx->scan_scratch.as<int>(), ctx->stream, &ctx->launches);
    pgt_launch_scatter_ids(ctx->cell_of.as<int>(), ctx->cell_fill.as<int>(), N, ctx->cell_count.as<int>(),
                           ctx->sorted_orig.as<int>(), ctx->stream, &ctx->launches);

    // live count and error flags to the host (mapped memory).  A frame of the same shape as the last one counted does not
    // wait for them: its arrays are sized from that frame's count with a margin, the kernels read the live count on the
    // device, and a count beyond the margin raises the sticky error word
    long long sig[8];
    frame_signature(ctx, sig);
    const bool defer = ctx->async_counts && ctx->known_nlive >= 0 && !memcmp(sig, ctx->known_sig, sizeof(sig));
    const long long bound = defer ? std::min<long long>(N, (long long)ctx->known_nlive + ctx->known_nlive / 8 + 1024) : N;
    pgt_launch_publish_count(ctx->cell_count.as<int>() + ncells, ctx->eflags.as<int>(), bound, ctx->eflags.as<int>() + 2, d_counts(ctx),
                             d_counts(ctx) + 4, ctx->stream, &ctx->launches);
    ctx->cells_deferred = defer;
    if (defer) {
        ctx->n_live = (int)bound;
        ctx->counts_pending = true;
    } else {
        CK(cudaStreamSynchronize(ctx->stream));
        ctx->counts_pending = false;
        const int* h = h_counts(ctx);
        if (h[4]) return sticky_error(ctx, h[4]);
        ctx->n_live = h[0];
        ctx->known_nlive = -1;                               // valid again once the bond stage has counted too
        memcpy(ctx->known_sig, sig, sizeof(sig));
    }

    const size_t nl = (size_t)std::max(ctx->n_live, 1);
    ENSURE(ctx->sx, sizeof(double) * nl);
    ENSURE(ctx->sy, sizeof(double) * nl);
    ENSURE(ctx->sxy, sizeof(double2) * nl);
    ENSURE(ctx->sd, sizeof(double) * nl);
    const bool kin = ctx->src_vel != nullptr;
    if (kin) {
        ENSURE(ctx->skxx, sizeof(double) * nl);
        ENSURE(ctx->skxy, sizeof(double) * nl);
        ENSURE(ctx->skyy, sizeof(double) * nl);
    }
    pgt_launch_cell_sort_gather(ctx->src_pos, ctx->src_diam, ctx->src_mass, ctx->src_vel, ctx->is_f32, ctx->box,
                                ctx->sorted_orig.as<int>(), ctx->cell_count.as<int>(), ncells, ctx->n_live, ctx->sx.as<double>(), ctx->sy.as<double>(),
                                ctx->sxy.as<double2>(), ctx->sd.as<double>(), kin ? ctx->skxx.as<double>() : nullptr,
                                kin ? ctx->skxy.as<double>() : nullptr, kin ? ctx->skyy.as<double>() : nullptr, ctx->stream, &ctx->launches);
    CK(cudaGetLastError());
    if (ctx->raw_cur >= 0) {                                 // stage 1 was the last reader of the raw frame
        CK(cudaEventRecord(ctx->ev_raw_free[ctx->raw_cur], ctx->stream));
        ctx->raw_free_valid[ctx->raw_cur] = true;
    }
    stage_end(ctx, PGT_STAGE_CELLS);
    ctx->cells_built = true;
    ctx->bonds_built = false;
    return PGT_OK;
}

int pgt_build_bonds(pgt_ctx* ctx)
{
    if (!ctx) return PGT_ERR_ARG;
    if (!ctx->cells_built) return fail(ctx, PGT_ERR_STATE, "pgt_build_bonds: cells not built");
    if (!ctx->have_pot) return fail(ctx, PGT_ERR_STATE, "pgt_build_bonds: no pair potential");
    CK(cudaSetDevice(ctx->device));
    if (ctx->dim == 3) {                                       // no records in 3-D: the gather searches the pairs itself
        ctx->n_hb = 0;
        ctx->bonds_built = true;
        return PGT_OK;
    }
    // deferred counts: the record arrays keep last frame's size plus a margin, the writer drops records beyond it and the
    // sticky word reports it; needs the single-search path and the warp-flattened writer (the one with the capacity guard)
    const bool defer = ctx->cells_deferred && ctx->counts_pending && ctx->known_nhb >= 0 && ctx->bond_single_search &&
                       ctx->bond_fill_flat && ctx->hb_pm_valid && ctx->n_live > 0 && ctx->n_live < (1 << 28);
    if (ctx->counts_pending && !defer) {                       // this stage counts exactly: so must the frame
        int rc = resolve_counts(ctx);
        if (rc != PGT_OK) return rc;
    }
    const int n = ctx->n_live;
    stage_begin(ctx, PGT_STAGE_BONDS);
    ENSURE(ctx->hb_cnt, sizeof(int) * ((size_t)n + 1));
    CK(cudaMemsetAsync(ctx->hb_cnt.as<int>() + n, 0, sizeof(int), ctx->stream));
    PgtFrameView fr = frame_view(ctx);
    // single search: count and keep the neighbours; a particle with more than PGT_NB_SLOTS bonds sends the frame
    // down the two-search path below
    bool listed = ctx->bond_single_search && n > 0 && n < (1 << 28);
    if (listed) {
        ENSURE(ctx->nb_list, sizeof(unsigned) * PGT_NB_SLOTS * (size_t)n);
        ENSURE(ctx->hb_pm, sizeof(unsigned short) * (size_t)n);
        CK(cudaMemsetAsync(ctx->eflags.p, 0, sizeof(int), ctx->stream));
        pgt_launch_bonds(4, ctx->pot, fr, ctx->box, ctx->cells, ctx->max_rcut, ctx->hb_cnt.as<int>(),
                         PgtBondOut{ ctx->hb_pm.as<double2>(), nullptr, ctx->eflags.as<int>() }, 1.0, ctx->nb_list.as<int>(), ctx->stream,
                         &ctx->launches);
    } else {
        pgt_launch_bonds(0, ctx->pot, fr, ctx->box, ctx->cells, ctx->max_rcut, ctx->hb_cnt.as<int>(), PgtBondOut{}, 1.0, nullptr,
                         ctx->stream, &ctx->launches);
    }
    pgt_launch_scan_exclusive(ctx->hb_cnt.as<int>(), ctx->hb_cnt.as<int>(), (long long)n + 1, ctx->scan_scratch.as<int>(),
                              ctx->stream, &ctx->launches);
    const long long bound = defer ? ctx->known_nhb + ctx->known_nhb / 8 + 4096 : 0x7FFFFFFFLL;
    pgt_launch_publish_count(ctx->hb_cnt.as<int>() + n, ctx->eflags.as<int>(), bound, ctx->eflags.as<int>() + 2, d_counts(ctx) + 2,
                             d_counts(ctx) + 4, ctx->stream, &ctx->launches);
    if (defer) {
        ctx->n_hb = ctx->known_nhb;                            // an estimate until the counts are resolved
    } else {
        CK(cudaStreamSynchronize(ctx->stream));
        const int* h = h_counts(ctx);
        if (h[2] < 0) return fail(ctx, PGT_ERR_DATA, "half-bond count overflows int32");
        if (listed && (h[3] & PGT_EFLAG_BOND_OVERFLOW)) {
            // the counts are right, the kept lists are truncated: fill by a second search
            listed = false;
            CK(cudaMemsetAsync(ctx->eflags.as<int>() + 2, 0, sizeof(int), ctx->stream));   // handled here, not an error
            h_counts(ctx)[4] = 0;
        } else if (h[4]) {
            return sticky_error(ctx, h[4]);
        }
        ctx->n_hb = h[2];
        ctx->known_nhb = ctx->n_hb;                            // the next frame of this shape may defer
        ctx->known_nlive = ctx->n_live;
    }
    ctx->hb_pm_valid = listed;
    const size_t nb = (size_t)std::max<long long>(defer ? bound : ctx->n_hb, 1);
    ENSURE(ctx->hb_b, sizeof(double2) * nb);
    ENSURE(ctx->hb_cf, sizeof(double2) * nb);
    ENSURE(ctx->hb_j, sizeof(int) * nb);
    ctx->hb_cap = (long long)nb;
    ctx->hb_cscale = ctx->have_cg ? 1.0 / ctx->cg.rcutsq : 1.0;
    fr = frame_view(ctx);
    const PgtBondOut out{ ctx->hb_b.as<double2>(), ctx->hb_cf.as<double2>(), ctx->hb_j.as<int>(), ctx->hb_cap };
    if (listed)
        pgt_launch_bonds(ctx->bond_fill_flat ? 5 : 6, ctx->pot, fr, ctx->box, ctx->cells, ctx->max_rcut, ctx->hb_cnt.as<int>(), out, ctx->hb_cscale,
                         ctx->nb_list.as<int>(), ctx->stream, &ctx->launches);
    else
        pgt_launch_bonds(1, ctx->pot, fr, ctx->box, ctx->cells, ctx->max_rcut, ctx->hb_cnt.as<int>(), out, ctx->hb_cscale, nullptr,
                         ctx->stream, &ctx->launches);
    CK(cudaGetLastError());
    stage_end(ctx, PGT_STAGE_BONDS);
    ctx->bonds_built = true;
    return PGT_OK;
}

int pgt_frame_counts(pgt_ctx* ctx, int64_t* n_live, int64_t* n_halfbonds, int64_t* n_cells)
{
    if (!ctx) return PGT_ERR_ARG;
    RESOLVE();
    if (n_live) *n_live = ctx->cells_built ? ctx->n_live : -1;
    if (n_halfbonds) *n_halfbonds = ctx->bonds_built ? ctx->n_hb : -1;
    if (n_cells) *n_cells = !ctx->cells_built ? -1 : (ctx->dim == 3 ? (int64_t)ctx->cells3.ncx * ctx->cells3.ncy * ctx->cells3.ncz : (int64_t)ctx->cells.ncx * ctx->cells.ncy);
    return PGT_OK;
}

// ---------------------------------------------------------------------------------- grid
int pgt_grid_set_cartesian(pgt_ctx* ctx, const double* xs, int nx, const double* ys, int ny)
{
    if (!ctx || !xs || !ys || nx < 0 || ny < 0) return fail(ctx, PGT_ERR_ARG, "pgt_grid_set_cartesian: bad argument");
    CK(cudaSetDevice(ctx->device));
    ENSURE(ctx->gxs, sizeof(double) * (size_t)std::max(nx, 1));
    ENSURE(ctx->gys, sizeof(double) * (size_t)std::max(ny, 1));
    CK(cudaMemcpyAsync(ctx->gxs.p, xs, sizeof(double) * nx, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(ctx->gys.p, ys, sizeof(double) * ny, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    // largest extent of an aligned tile of grid points (sizes the quad gather's shared-memory rectangle)
    auto span = [](const double* v, int n, int t) {
        double s = 0.0;
        for (int a = 0; a < n; a += t) {
            double lo = v[a], hi = v[a];
            for (int k = a + 1; k < std::min(a + t, n); ++k) { lo = std::min(lo, v[k]); hi = std::max(hi, v[k]); }
            s = std::max(s, hi - lo);
        }
        return s;
    };
    ctx->grid_monotone = std::is_sorted(xs, xs + nx) && std::is_sorted(ys, ys + ny);
    ctx->h_xs.assign(xs, xs + nx);
    for (int w = 1; w <= 4; ++w) {
        ctx->span_x[w - 1] = nx > 0 ? span(xs, nx, 8 * w) : -1.0;
        ctx->span_y[w - 1] = ny > 0 ? span(ys, ny, 4 * w) : -1.0;
    }
    ctx->grid_mode = 0;
    ctx->nx = nx;
    ctx->ny = ny;
    ctx->G = (long long)nx * ny;
    ctx->g0 = 0;
    ctx->g1 = ctx->G;
    return PGT_OK;
}

int pgt_grid_set_points(pgt_ctx* ctx, const double* xyz, int64_t G)
{
    if (!ctx || (!xyz && G > 0) || G < 0) return fail(ctx, PGT_ERR_ARG, "pgt_grid_set_points: bad argument");
    CK(cudaSetDevice(ctx->device));
    ENSURE(ctx->gpts, sizeof(double) * 3 * (size_t)std::max<long long>(G, 1));
    if (G > 0) CK(cudaMemcpyAsync(ctx->gpts.p, xyz, sizeof(double) * 3 * (size_t)G, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    ctx->grid_mode = 1;
    ctx->G = G;
    ctx->g0 = 0;
    ctx->g1 = G;
    return PGT_OK;
}

int pgt_grid_set_shard(pgt_ctx* ctx, int64_t g0, int64_t g1)
{
    if (!ctx) return PGT_ERR_ARG;
    if (ctx->grid_mode < 0) return fail(ctx, PGT_ERR_STATE, "pgt_grid_set_shard: no grid");
    if (g0 < 0 || g1 < g0 || g1 > ctx->G) return fail(ctx, PGT_ERR_ARG, "shard [%lld,%lld) outside [0,%lld)", (long long)g0, (long long)g1, ctx->G);
    ctx->g0 = g0;
    ctx->g1 = g1;
    return PGT_OK;
}

// ---------------------------------------------------------------------------------- compute
static PgtIKArgs ik_args_impl(pgt_ctx* ctx, unsigned obs_mask);
static int ensure_disp(pgt_ctx* ctx);
#define ik_args ik_args_impl
int pgt_ik_compute(pgt_ctx* ctx, unsigned obs_mask)
{
    if (!ctx) return PGT_ERR_ARG;
    if (!ctx->cells_built) return fail(ctx, PGT_ERR_STATE, "pgt_ik_compute: cells not built");
    if ((obs_mask & PGT_OBS_VIRIAL) && !ctx->bonds_built) return fail(ctx, PGT_ERR_STATE, "pgt_ik_compute: bonds not built");
    if (!ctx->have_cg) return fail(ctx, PGT_ERR_STATE, "pgt_ik_compute: no coarse-graining function");
    if (ctx->grid_mode < 0) return fail(ctx, PGT_ERR_STATE, "pgt_ik_compute: no grid");
    if (!(obs_mask & (PGT_OBS_VIRIAL | PGT_OBS_KINETIC | PGT_OBS_DENSITY | PGT_OBS_ELASTIC))) return fail(ctx, PGT_ERR_ARG, "pgt_ik_compute: empty observable mask");
    if ((obs_mask & PGT_OBS_ELASTIC) && !ctx->have_pot) return fail(ctx, PGT_ERR_STATE, "pgt_ik_compute: no pair potential");
    if (ctx->dim == 3) return ik_compute3(ctx, obs_mask);
    if (ctx->cg.rcut >= ctx->box.L[0] || ctx->cg.rcut >= ctx->box.L[1])
        return fail(ctx, PGT_ERR_DATA, "cg_rcut %g is not smaller than the box", ctx->cg.rcut);
    if (ctx->n_live >= (1 << 28))
        return fail(ctx, PGT_ERR_UNSUPPORTED, "pgt_ik_compute: %d live particles; the gather's packed lists address 2^28", ctx->n_live);
    if (ctx->quad_mode == PGT_QUAD_ADAPTIVE || (obs_mask & PGT_OBS_ELASTIC)) RESOLVE();   // not throughput paths: exact counts
    if (ctx->culled) {
        double lo = 0.0, hi = 0.0;
        if (!shard_slab(ctx, ctx->cells, &lo, &hi) || lo < ctx->slab_lo || hi > ctx->slab_hi)
            return fail(ctx, PGT_ERR_STATE, "pgt_ik_compute: the frame was built for another grid shard (rebuild the cells after changing the grid, the shard or the coarse-graining radius)");
    }
    CK(cudaSetDevice(ctx->device));
    const long long gl = ctx->g1 - ctx->g0;
    ctx->field_g0 = ctx->g0;                                 // read-backs, device_ptr and the NCCL gather size their copies from these
    ctx->field_g1 = ctx->g1;
    const size_t n = (size_t)std::max<long long>(gl, 1);
    if (obs_mask & PGT_OBS_VIRIAL) ENSURE(ctx->out_virial, sizeof(double) * 4 * n);
    if (obs_mask & PGT_OBS_KINETIC) ENSURE(ctx->out_kinetic, sizeof(double) * 4 * n);
    if (obs_mask & PGT_OBS_DENSITY) ENSURE(ctx->out_density, sizeof(double) * n);
    if (obs_mask & PGT_OBS_ELASTIC) {
        ENSURE(ctx->out_elastic, sizeof(double) * 4 * n);
        const int rc = ensure_disp(ctx);
        if (rc != PGT_OK) return rc;
    }

    if (ctx->dl_pending) CK(cudaStreamWaitEvent(ctx->stream, ctx->ev_dl_done, 0));   // the field buffers are still being read back
    stage_begin(ctx, PGT_STAGE_IK);
    if ((obs_mask & PGT_OBS_VIRIAL) && ctx->hb_cscale != 1.0 / ctx->cg.rcutsq) {
        // the CG radius changed after the bond list was built: refresh the records' scaled |b|^2
        RESOLVE();                                           // the exact record count
        ctx->hb_cscale = 1.0 / ctx->cg.rcutsq;
        pgt_launch_rescale_bonds(ctx->hb_b.as<double2>(), ctx->hb_cf.as<double2>(), ctx->n_hb, ctx->hb_cscale, ctx->stream, &ctx->launches);
    }
    const unsigned mask_all = obs_mask;
    obs_mask &= ~PGT_OBS_ELASTIC;
    PgtIKArgs a = ik_args(ctx, obs_mask);
    bool check_qag = false;
    if (gl > 0 && (mask_all & PGT_OBS_ELASTIC)) {
        const bool adaptive = ctx->quad_mode == PGT_QUAD_ADAPTIVE;
        if (adaptive) { CK(cudaMemsetAsync(ctx->eflags.p, 0, sizeof(int), ctx->stream)); check_qag = true; }
        CK(pgt_launch_ik_elastic(ctx->pot, a, ctx->disp.as<double>(), ctx->max_rcut, adaptive ? 1 : 0, ctx->out_elastic.as<double>(),
                                 ctx->eflags.as<int>(), ctx->stream, &ctx->launches));
    }
    if (obs_mask == 0) {
    } else if (ctx->quad_mode == PGT_QUAD_ADAPTIVE && (obs_mask & PGT_OBS_VIRIAL)) {
        // local observables only need Delta(): the regular kernel; the virial field: the QAG kernel
        if (gl > 0 && (obs_mask & ~PGT_OBS_VIRIAL)) {
            PgtIKArgs al = a;
            al.obs_mask = obs_mask & ~PGT_OBS_VIRIAL;
            CK(pgt_launch_ik(al, 2, ctx->stream, &ctx->launches));
        }
        if (gl > 0) {
            if (!check_qag) CK(cudaMemsetAsync(ctx->eflags.p, 0, sizeof(int), ctx->stream));
            CK(pgt_launch_ik_adaptive(a, ctx->eflags.as<int>(), ctx->stream, &ctx->launches));
            check_qag = true;
        }
    } else if (gl > 0) CK(pgt_launch_ik(a, ctx->ik_variant, ctx->stream, &ctx->launches));
    stage_end(ctx, PGT_STAGE_IK);
    if (check_qag) {
        pgt_launch_publish(ctx->eflags.p, 1, nullptr, 0, ctx->d_small, ctx->stream, &ctx->launches);
        CK(cudaStreamSynchronize(ctx->stream));
        const int fl = reinterpret_cast<int*>(ctx->h_small)[0];
        if (fl & PGT_EFLAG_QAG_CAPACITY)
            return fail(ctx, PGT_ERR_UNSUPPORTED, "adaptive bond quadrature: a bond needs more than %d intervals on device", PGT_QAG_CAP);
        if (fl & PGT_EFLAG_QAG_FAILED)
            return fail(ctx, PGT_ERR_DATA, "adaptive bond quadrature did not reach the tolerance within maxiter = %d intervals "
                        "(GSL's error handler aborts the reference here)", ctx->qag_limit);
    }
    ctx->last_ik_mask = mask_all;
    return PGT_OK;
}

static PgtIKArgs ik_args_impl(pgt_ctx* ctx, unsigned obs_mask)
{
    PgtIKArgs a{};
    a.fr = frame_view(ctx);
    a.grid.mode = ctx->grid_mode;
    a.grid.nx = ctx->nx;
    a.grid.ny = ctx->ny;
    a.grid.xs = ctx->gxs.as<double>();
    a.grid.ys = ctx->gys.as<double>();
    a.grid.pts = ctx->gpts.as<double>();
    a.grid.g0 = ctx->g0;
    a.grid.g1 = ctx->g1;
    a.grid.monotone = ctx->grid_monotone ? 1 : 0;
    for (int w = 0; w < 4; ++w) { a.grid.span_x[w] = ctx->span_x[w]; a.grid.span_y[w] = ctx->span_y[w]; }
    a.box = ctx->box;
    a.cells = ctx->cells;
    a.cg = ctx->cg;
    a.inv_rcutsq = 1.0 / ctx->cg.rcutsq;
    a.n_frame = ctx->N;
    a.max_rcut = ctx->max_rcut;
    a.nq = ctx->nq;
    a.qag_epsabs = ctx->qag_epsabs;
    a.qag_epsrel = ctx->qag_epsrel;
    a.qag_limit = ctx->qag_limit;
    a.gl5_h1 = ctx->gl5[0]; a.gl5_h1sq = ctx->gl5[1]; a.gl5_h2 = ctx->gl5[2]; a.gl5_h2sq = ctx->gl5[3];
    a.gl5_w0 = ctx->gl5[4]; a.gl5_w1 = ctx->gl5[5]; a.gl5_w2 = ctx->gl5[6];
    for (int k = 0; k < 5; ++k) { a.gl_s[k] = ctx->gl_s[k]; a.gl_w[k] = ctx->gl_w[k]; }
    a.gl_nodes = ctx->gl_nodes.as<double>();
    a.gl_weights = ctx->gl_weights.as<double>();
    a.out_virial = ctx->out_virial.as<double>();
    a.out_kinetic = ctx->out_kinetic.as<double>();
    a.out_density = ctx->out_density.as<double>();
    a.obs_mask = obs_mask;
    return a;
}

int pgt_ik_count_work(pgt_ctx* ctx, int64_t* sum_unique_bonds, int64_t* sum_halfbond_items, int64_t* sum_particles)
{
    if (ctx) RESOLVE();
    if (!ctx) return PGT_ERR_ARG;
    if (ctx->dim != 2) return fail(ctx, PGT_ERR_UNSUPPORTED, "pgt_ik_count_work: 2-D frames only");
    if (!ctx->bonds_built || !ctx->have_cg || ctx->grid_mode < 0) return fail(ctx, PGT_ERR_STATE, "pgt_ik_count_work: frame, bonds, cg and grid required");
    CK(cudaSetDevice(ctx->device));
    DevBuf out;
    ENSURE(out, sizeof(unsigned long long) * 4);
    CK(cudaMemsetAsync(out.p, 0, sizeof(unsigned long long) * 4, ctx->stream));
    PgtIKArgs a = ik_args_impl(ctx, PGT_OBS_VIRIAL);
    pgt_launch_ik_count(a, out.as<unsigned long long>(), ctx->stream, &ctx->launches);
    unsigned long long h[3] = { 0, 0, 0 };
    CK(cudaMemcpyAsync(h, out.p, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    cudaFree(out.p);
    if (sum_unique_bonds) *sum_unique_bonds = (int64_t)h[0];
    if (sum_halfbond_items) *sum_halfbond_items = (int64_t)h[1];
    if (sum_particles) *sum_particles = (int64_t)h[2];
    return PGT_OK;
}

int pgt_event_record(pgt_ctx* ctx, int slot)
{
    if (!ctx || slot < 0 || slot >= 16) return fail(ctx, PGT_ERR_ARG, "pgt_event_record: slot out of range");
    CK(cudaSetDevice(ctx->device));
    if (!ctx->user_ev[slot]) CK(cudaEventCreate(&ctx->user_ev[slot]));
    CK(cudaEventRecord(ctx->user_ev[slot], ctx->stream));
    return PGT_OK;
}

int pgt_event_elapsed_ms(pgt_ctx* ctx, int a, int b, float* ms)
{
    if (!ctx || !ms || a < 0 || a >= 16 || b < 0 || b >= 16 || !ctx->user_ev[a] || !ctx->user_ev[b])
        return fail(ctx, PGT_ERR_ARG, "pgt_event_elapsed_ms: bad slots");
    CK(cudaEventSynchronize(ctx->user_ev[b]));
    CK(cudaEventElapsedTime(ms, ctx->user_ev[a], ctx->user_ev[b]));
    return PGT_OK;
}

int pgt_host_alloc(void** ptr, uint64_t bytes)
{
    if (!ptr) return PGT_ERR_ARG;
    cudaError_t e = cudaHostAlloc(ptr, bytes ? bytes : 1, cudaHostAllocDefault);
    if (e != cudaSuccess) { g_create_error = cudaGetErrorString(e); *ptr = nullptr; return PGT_ERR_CUDA; }
    return PGT_OK;
}

int pgt_host_register(void* ptr, uint64_t bytes)
{
    if (!ptr || !bytes) return PGT_ERR_ARG;
    const cudaError_t e = cudaHostRegister(ptr, bytes, cudaHostRegisterDefault);
    if (e != cudaSuccess) { cudaGetLastError(); return PGT_ERR_CUDA; }
    return PGT_OK;
}

int pgt_host_unregister(void* ptr)
{
    if (!ptr) return PGT_ERR_ARG;
    if (cudaHostUnregister(ptr) != cudaSuccess) { cudaGetLastError(); return PGT_ERR_CUDA; }
    return PGT_OK;
}

int pgt_host_free(void* ptr)
{
    if (ptr && cudaFreeHost(ptr) != cudaSuccess) return PGT_ERR_CUDA;
    return PGT_OK;
}

int pgt_set_option(pgt_ctx* ctx, const char* name, double value)
{
    if (!ctx || !name) return PGT_ERR_ARG;
    if (!strcmp(name, "ik_variant")) {
        const int l = (int)value;
        if (l != 0 && l != 2 && l != 7) return fail(ctx, PGT_ERR_ARG, "ik_variant must be 0 (auto), 2 (two-phase) or 7 (quad)");
        ctx->ik_variant = l;
        return PGT_OK;
    }
    if (!strcmp(name, "async_counts")) {
        ctx->async_counts = value != 0.0;
        ctx->known_nlive = -1;
        ctx->known_nhb = -1;
        return PGT_OK;
    }
    if (!strcmp(name, "slab_cull")) {
        ctx->slab_cull = value != 0.0;
        ctx->cells_built = false;
        ctx->bonds_built = false;
        return PGT_OK;
    }
    if (!strcmp(name, "bond_fill_flat")) {
        ctx->bond_fill_flat = value != 0.0;
        ctx->bonds_built = false;
        return PGT_OK;
    }
    if (!strcmp(name, "bond_single_search")) {
        ctx->bond_single_search = value != 0.0;
        ctx->bonds_built = false;
        return PGT_OK;
    }
    return fail(ctx, PGT_ERR_ARG, "unknown option %s", name);
}

static int field_info(pgt_ctx* ctx, int field_id, DevBuf** buf, int* ncomp)
{
    switch (field_id) {
    case PGT_FIELD_VIRIAL: *buf = &ctx->out_virial; *ncomp = ctx->dim * ctx->dim; return (ctx->last_ik_mask & PGT_OBS_VIRIAL) ? 0 : -1;
    case PGT_FIELD_KINETIC: *buf = &ctx->out_kinetic; *ncomp = ctx->dim * ctx->dim; return (ctx->last_ik_mask & PGT_OBS_KINETIC) ? 0 : -1;
    case PGT_FIELD_DENSITY: *buf = &ctx->out_density; *ncomp = 1; return (ctx->last_ik_mask & PGT_OBS_DENSITY) ? 0 : -1;
    case PGT_FIELD_ELASTIC: *buf = &ctx->out_elastic; *ncomp = ctx->dim * ctx->dim; return (ctx->last_ik_mask & PGT_OBS_ELASTIC) ? 0 : -1;
    }
    return -2;
}

static int ik_download_impl(pgt_ctx* ctx, int field_id, double* out, bool wait);
int pgt_ik_download(pgt_ctx* ctx, int field_id, double* out) { return ik_download_impl(ctx, field_id, out, true); }
int pgt_ik_download_async(pgt_ctx* ctx, int field_id, double* out) { return ik_download_impl(ctx, field_id, out, false); }

static int ik_download_impl(pgt_ctx* ctx, int field_id, double* out, bool wait)
{
    if (!ctx || !out) return fail(ctx, PGT_ERR_ARG, "pgt_ik_download: null argument");
    DevBuf* b;
    int nc;
    int rc = field_info(ctx, field_id, &b, &nc);
    if (rc == -2) return fail(ctx, PGT_ERR_ARG, "unknown field %d", field_id);
    if (rc == -1) return fail(ctx, PGT_ERR_STATE, "field %d was not computed by the last pgt_ik_compute", field_id);
    CK(cudaSetDevice(ctx->device));
    const long long gl = ctx->field_g1 - ctx->field_g0;   // the shard the field was computed for
    if (wait) {
        stage_begin(ctx, PGT_STAGE_DOWNLOAD);
        if (gl > 0) CK(cudaMemcpyAsync(out, b->p, sizeof(double) * nc * (size_t)gl, cudaMemcpyDeviceToHost, ctx->stream));
        stage_end(ctx, PGT_STAGE_DOWNLOAD);
        CK(cudaStreamSynchronize(ctx->stream));
        return resolve_counts(ctx);
    }
    // asynchronous: the copy runs on its own stream behind the gather, so that the next frame's upload and
    // stages (main stream) are not queued behind it
    CK(cudaEventRecord(ctx->ev_field_ready, ctx->stream));
    CK(cudaStreamWaitEvent(ctx->dl_stream, ctx->ev_field_ready, 0));
    if (gl > 0) CK(cudaMemcpyAsync(out, b->p, sizeof(double) * nc * (size_t)gl, cudaMemcpyDeviceToHost, ctx->dl_stream));
    CK(cudaEventRecord(ctx->ev_dl_done, ctx->dl_stream));
    CK(cudaEventRecord(ctx->ev_dl_copy[ctx->dl_seq & 3], ctx->dl_stream));
    ctx->dl_seq++;
    ctx->dl_pending = true;
    return PGT_OK;
}

int pgt_ik_download_sym_async(pgt_ctx* ctx, int field_id, double* out)
{
    if (!ctx || !out) return fail(ctx, PGT_ERR_ARG, "pgt_ik_download_sym_async: null argument");
    if (ctx->dim != 2) return fail(ctx, PGT_ERR_UNSUPPORTED, "pgt_ik_download_sym_async: 2-D fields only (read a 3-D field with pgt_ik_download)");
    DevBuf* b;
    int nc;
    int rc = field_info(ctx, field_id, &b, &nc);
    if (rc == -2) return fail(ctx, PGT_ERR_ARG, "unknown field %d", field_id);
    if (rc == -1) return fail(ctx, PGT_ERR_STATE, "field %d was not computed by the last pgt_ik_compute", field_id);
    if (nc != 4) return fail(ctx, PGT_ERR_ARG, "field %d is not a 2x2 tensor field", field_id);
    CK(cudaSetDevice(ctx->device));
    const long long gl = ctx->field_g1 - ctx->field_g0;   // the shard the field was computed for
    // the pack buffer may still be read back by the previous copy: the read-back stream is in order, the pack below queues
    // behind it (only a re-allocation of the buffer has to wait on the host)
    if (ctx->dl_pending && sizeof(double) * 3 * (size_t)std::max<long long>(gl, 1) > ctx->pack.cap) CK(cudaStreamSynchronize(ctx->dl_stream));
    ENSURE(ctx->pack, sizeof(double) * 3 * (size_t)std::max<long long>(gl, 1));
    CK(cudaEventRecord(ctx->ev_field_ready, ctx->stream));
    CK(cudaStreamWaitEvent(ctx->dl_stream, ctx->ev_field_ready, 0));
    pgt_launch_pack_sym(static_cast<const double*>(b->p), ctx->pack.as<double>(), gl, ctx->dl_stream, &ctx->launches);
    // the next gather may overwrite the field as soon as it has been packed; the copy reads the pack buffer
    CK(cudaEventRecord(ctx->ev_dl_done, ctx->dl_stream));
    if (gl > 0) CK(cudaMemcpyAsync(out, ctx->pack.p, sizeof(double) * 3 * (size_t)gl, cudaMemcpyDeviceToHost, ctx->dl_stream));
    CK(cudaEventRecord(ctx->ev_dl_copy[ctx->dl_seq & 3], ctx->dl_stream));
    ctx->dl_seq++;
    ctx->dl_pending = true;
    return PGT_OK;
}

int pgt_ik_download_wait(pgt_ctx* ctx)
{
    if (!ctx) return PGT_ERR_ARG;
    if (ctx->dl_pending) {
        CK(cudaStreamSynchronize(ctx->dl_stream));
        ctx->dl_pending = false;
        ctx->dl_waited = ctx->dl_seq;
        // the frame whose field has just landed published its counts before its gather ran: an error word raised by it
        // (or by a later frame) is visible now, without waiting for the frames still in flight on the main stream
        if (ctx->counts_pending && h_counts(ctx)[4]) {
            CK(cudaStreamSynchronize(ctx->stream));
            ctx->counts_pending = false;
            return sticky_error(ctx, h_counts(ctx)[4]);
        }
    }
    return PGT_OK;
}

int pgt_ik_download_wait_keep(pgt_ctx* ctx, int keep)
{
    if (!ctx) return PGT_ERR_ARG;
    if (keep < 0 || keep > 3) return fail(ctx, PGT_ERR_ARG, "pgt_ik_download_wait_keep: keep must be 0..3");
    const long long upto = ctx->dl_seq - keep;               // read-backs [0, upto) must have landed
    if (upto > ctx->dl_waited) {
        CK(cudaEventSynchronize(ctx->ev_dl_copy[(upto - 1) & 3]));   // the stream is in order: the youngest of them suffices
        ctx->dl_waited = upto;
        if (ctx->counts_pending && h_counts(ctx)[4]) {       // see pgt_ik_download_wait
            CK(cudaStreamSynchronize(ctx->stream));
            ctx->counts_pending = false;
            return sticky_error(ctx, h_counts(ctx)[4]);
        }
    }
    return PGT_OK;
}

int pgt_ik_device_ptr(pgt_ctx* ctx, int field_id, void** d_ptr, int64_t* n_doubles)
{
    if (!ctx || !d_ptr) return fail(ctx, PGT_ERR_ARG, "pgt_ik_device_ptr: null argument");
    DevBuf* b;
    int nc;
    int rc = field_info(ctx, field_id, &b, &nc);
    if (rc == -2) return fail(ctx, PGT_ERR_ARG, "unknown field %d", field_id);
    if (rc == -1) return fail(ctx, PGT_ERR_STATE, "field %d was not computed by the last pgt_ik_compute", field_id);
    *d_ptr = b->p;
    if (n_doubles) *n_doubles = (int64_t)nc * (ctx->field_g1 - ctx->field_g0);
    return PGT_OK;
}

int pgt_thermo_compute(pgt_ctx* ctx, unsigned obs_mask, double* out)
{
    if (ctx) RESOLVE();
    if (!ctx || !out) return fail(ctx, PGT_ERR_ARG, "pgt_thermo_compute: null argument");
    if (!ctx->cells_built) return fail(ctx, PGT_ERR_STATE, "pgt_thermo_compute: cells not built");
    if (!ctx->have_pot) return fail(ctx, PGT_ERR_STATE, "pgt_thermo_compute: no pair potential");
    if (ctx->culled) return fail(ctx, PGT_ERR_STATE, "pgt_thermo_compute: the frame holds a grid shard's particles only (set the whole grid or option slab_cull = 0 and rebuild the cells)");
    if (ctx->dim == 3) return thermo3(ctx, obs_mask, false, out);
    CK(cudaSetDevice(ctx->device));
    const int blocks = pgt_thermo_blocks(ctx->n_live);
    const int p_begin = (int)((long long)ctx->n_live * ctx->thermo_part / ctx->thermo_nparts);
    const int p_end = (int)((long long)ctx->n_live * (ctx->thermo_part + 1) / ctx->thermo_nparts);
    ENSURE(ctx->thermo_partials, sizeof(double) * 25 * (size_t)blocks);
    ENSURE(ctx->thermo_out, sizeof(double) * 32);
    stage_begin(ctx, PGT_STAGE_THERMO);
    pgt_launch_thermo(ctx->pot, frame_view(ctx), ctx->box, ctx->cells, ctx->max_rcut, obs_mask, 1 / ctx->vol,
                      ctx->thermo_partials.as<double>(), blocks, ctx->thermo_out.as<double>(), p_begin, p_end, ctx->stream, &ctx->launches);
    CK(cudaGetLastError());
    stage_end(ctx, PGT_STAGE_THERMO);
    pgt_launch_publish(ctx->thermo_out.p, 50, nullptr, 0, ctx->d_small, ctx->stream, &ctx->launches);
    CK(cudaStreamSynchronize(ctx->stream));
    const double* h = ctx->h_small;
    // device order: virial 4 | kinetic 4 | energy | born 16  ->  API order (same, documented in the header)
    for (int k = 0; k < 25; ++k) out[k] = h[k];
    return PGT_OK;
}

// The same sums without a host synchronisation: the 25 doubles are copied into `out` (page-locked memory for a copy that
// really is asynchronous) behind the kernels on the context's stream, valid after pgt_synchronize.  With deferred counts
// (option async_counts) a trajectory loop of upload / cells / thermo never waits for the device.
int pgt_thermo_compute_async(pgt_ctx* ctx, unsigned obs_mask, double* out)
{
    if (!ctx || !out) return fail(ctx, PGT_ERR_ARG, "pgt_thermo_compute_async: null argument");
    if (!ctx->cells_built) return fail(ctx, PGT_ERR_STATE, "pgt_thermo_compute_async: cells not built");
    if (!ctx->have_pot) return fail(ctx, PGT_ERR_STATE, "pgt_thermo_compute_async: no pair potential");
    if (ctx->culled) return fail(ctx, PGT_ERR_STATE, "pgt_thermo_compute_async: the frame holds a grid shard's particles only");
    if (ctx->dim != 2) return fail(ctx, PGT_ERR_UNSUPPORTED, "pgt_thermo_compute_async: 2-D frames only");
    if (ctx->thermo_nparts > 1) RESOLVE();                     // the particle shares are cut from the exact live count
    CK(cudaSetDevice(ctx->device));
    const int blocks = pgt_thermo_blocks(ctx->n_live);
    const int p_begin = (int)((long long)ctx->n_live * ctx->thermo_part / ctx->thermo_nparts);
    const int p_end = (int)((long long)ctx->n_live * (ctx->thermo_part + 1) / ctx->thermo_nparts);
    ENSURE(ctx->thermo_partials, sizeof(double) * 25 * (size_t)blocks);
    ENSURE(ctx->thermo_out, sizeof(double) * 32);
    stage_begin(ctx, PGT_STAGE_THERMO);
    pgt_launch_thermo(ctx->pot, frame_view(ctx), ctx->box, ctx->cells, ctx->max_rcut, obs_mask, 1 / ctx->vol,
                      ctx->thermo_partials.as<double>(), blocks, ctx->thermo_out.as<double>(), p_begin, p_end, ctx->stream, &ctx->launches);
    CK(cudaGetLastError());
    stage_end(ctx, PGT_STAGE_THERMO);
    CK(cudaMemcpyAsync(out, ctx->thermo_out.p, sizeof(double) * 25, cudaMemcpyDeviceToHost, ctx->stream));
    return PGT_OK;
}

// ---- pgt_thermo_compute_static: the thermo pair sums with no allocation at all -------------------------------------
// The per-block partial sums and the reduced result live in ONE file-scope device array instead of the context's
// cudaMalloc'ed buffers.  Every device that loads this library holds the array until the process exits; every context
// and every call on a device shares it (calls that overlap on the GPU race) and it keeps the last call's contents.
#define PGT_STATIC_SCRATCH_DOUBLES (PGT_THERMO_STATIC_MAX_BLOCKS * 25 + 32)     /* partials | result slot (25 used)  */
static __device__ __align__(256) double g_thermo_static_scratch[PGT_STATIC_SCRATCH_DOUBLES];

// Device code of this file names the array directly: the result slot starts every call at zero.
static __global__ void __launch_bounds__(32) k_thermo_static_clear()
{
    g_thermo_static_scratch[PGT_THERMO_STATIC_MAX_BLOCKS * 25 + threadIdx.x] = 0.0;
}

namespace {
// In host code `g_thermo_static_scratch` is a host-side stand-in, not a device address: the address each device holds the
// array at comes from cudaGetSymbolAddress, asked once per device with that device current.
constexpr int STATIC_MAX_DEVICES = 64;
double* g_static_ptr[STATIC_MAX_DEVICES] = {};
pthread_mutex_t g_static_mu = PTHREAD_MUTEX_INITIALIZER;

int static_scratch_ptr(pgt_ctx* ctx, double** out)
{
    if (ctx->device < 0 || ctx->device >= STATIC_MAX_DEVICES) return fail(ctx, PGT_ERR_ARG, "pgt_thermo_compute_static: device %d beyond %d", ctx->device, STATIC_MAX_DEVICES);
    pthread_mutex_lock(&g_static_mu);
    double* p = g_static_ptr[ctx->device];
    cudaError_t e = cudaSuccess;
    if (!p) {
        e = cudaSetDevice(ctx->device);
        if (e == cudaSuccess) e = cudaGetSymbolAddress((void**)&p, g_thermo_static_scratch);
        if (e == cudaSuccess) g_static_ptr[ctx->device] = p;
    }
    pthread_mutex_unlock(&g_static_mu);
    if (e != cudaSuccess || !p) return fail(ctx, PGT_ERR_CUDA, "cudaGetSymbolAddress(g_thermo_static_scratch) on device %d failed: %s", ctx->device, cudaGetErrorString(e));
    *out = p;
    return PGT_OK;
}
}  // namespace

int pgt_thermo_compute_static(pgt_ctx* ctx, unsigned obs_mask, double* out)
{
    if (ctx) RESOLVE();
    if (!ctx || !out) return fail(ctx, PGT_ERR_ARG, "pgt_thermo_compute_static: null argument");
    if (!ctx->cells_built) return fail(ctx, PGT_ERR_STATE, "pgt_thermo_compute_static: cells not built");
    if (!ctx->have_pot) return fail(ctx, PGT_ERR_STATE, "pgt_thermo_compute_static: no pair potential");
    if (ctx->culled) return fail(ctx, PGT_ERR_STATE, "pgt_thermo_compute_static: the frame holds a grid shard's particles only (set the whole grid or option slab_cull = 0 and rebuild the cells)");
    if (ctx->dim != 2) return fail(ctx, PGT_ERR_UNSUPPORTED, "pgt_thermo_compute_static: 2-D frames only");
    if (ctx->n_live > PGT_THERMO_STATIC_MAX_PARTICLES)
        return fail(ctx, PGT_ERR_ARG, "pgt_thermo_compute_static: the frame has %lld live particles, the static scratch serves at most %d (use pgt_thermo_compute)",
                    (long long)ctx->n_live, PGT_THERMO_STATIC_MAX_PARTICLES);
    const int blocks = pgt_thermo_blocks(ctx->n_live);
    if (blocks > PGT_THERMO_STATIC_MAX_BLOCKS)
        return fail(ctx, PGT_ERR_ARG, "pgt_thermo_compute_static: %d blocks of partial sums, the static scratch holds %d", blocks, PGT_THERMO_STATIC_MAX_BLOCKS);
    CK(cudaSetDevice(ctx->device));
    double* scratch = nullptr;
    int rc = static_scratch_ptr(ctx, &scratch);
    if (rc != PGT_OK) return rc;
    double* result = scratch + (size_t)PGT_THERMO_STATIC_MAX_BLOCKS * 25;
    const int p_begin = (int)((long long)ctx->n_live * ctx->thermo_part / ctx->thermo_nparts);
    const int p_end = (int)((long long)ctx->n_live * (ctx->thermo_part + 1) / ctx->thermo_nparts);
    k_thermo_static_clear<<<1, 32, 0, ctx->stream>>>();
    ctx->launches.n++;
    // same kernels, same block partition as pgt_thermo_compute: every block writes all 25 of its partial sums, so the
    // partials need no clearing
    pgt_launch_thermo(ctx->pot, frame_view(ctx), ctx->box, ctx->cells, ctx->max_rcut, obs_mask, 1 / ctx->vol, scratch, blocks, result,
                      p_begin, p_end, ctx->stream, &ctx->launches);
    CK(cudaGetLastError());
    pgt_launch_publish(result, 50, nullptr, 0, ctx->d_small, ctx->stream, &ctx->launches);
    CK(cudaStreamSynchronize(ctx->stream));
    const double* h = ctx->h_small;
    for (int k = 0; k < 25; ++k) out[k] = h[k];
    return PGT_OK;
}

static int ensure_disp(pgt_ctx* ctx)
{
    const size_t bytes = sizeof(double) * 3 * (size_t)std::max<long long>(ctx->N, 1);
    if (!ctx->have_disp) {                                   // ParticleSystem.cc:28: displacement = 0 until set
        ENSURE(ctx->disp, bytes);
        CK(cudaMemsetAsync(ctx->disp.p, 0, bytes, ctx->stream));
        ctx->have_disp = true;
    }
    return PGT_OK;
}

int pgt_frame_set_displacement(pgt_ctx* ctx, const double* disp_xyz, int64_t N)
{
    if (!ctx || (!disp_xyz && N > 0)) return fail(ctx, PGT_ERR_ARG, "pgt_frame_set_displacement: null argument");
    if (N != ctx->N) return fail(ctx, PGT_ERR_ARG, "displacement has %lld rows, the frame has %lld particles", (long long)N, (long long)ctx->N);
    CK(cudaSetDevice(ctx->device));
    const size_t bytes = sizeof(double) * 3 * (size_t)std::max<long long>(N, 1);
    ENSURE(ctx->disp, bytes);
    if (N > 0) CK(cudaMemcpyAsync(ctx->disp.p, disp_xyz, sizeof(double) * 3 * (size_t)N, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));                  // the caller's buffer may be pageable and short-lived
    ctx->have_disp = true;
    return PGT_OK;
}

int pgt_thermo_elastic(pgt_ctx* ctx, double* out)
{
    if (ctx) RESOLVE();
    if (!ctx || !out) return fail(ctx, PGT_ERR_ARG, "pgt_thermo_elastic: null argument");
    if (!ctx->cells_built) return fail(ctx, PGT_ERR_STATE, "pgt_thermo_elastic: cells not built");
    if (!ctx->have_pot) return fail(ctx, PGT_ERR_STATE, "pgt_thermo_elastic: no pair potential");
    if (ctx->dim == 3) return thermo3(ctx, 0, true, out);
    CK(cudaSetDevice(ctx->device));
    int rc = ensure_disp(ctx);
    if (rc != PGT_OK) return rc;
    const int blocks = pgt_thermo_blocks(ctx->n_live);
    ENSURE(ctx->thermo_partials, sizeof(double) * 25 * (size_t)blocks);
    ENSURE(ctx->thermo_out, sizeof(double) * 32);
    stage_begin(ctx, PGT_STAGE_THERMO);
    pgt_launch_thermo_elastic(ctx->pot, frame_view(ctx), ctx->box, ctx->cells, ctx->max_rcut, ctx->disp.as<double>(), 1 / ctx->vol,
                              ctx->thermo_partials.as<double>(), blocks, ctx->thermo_out.as<double>(), ctx->stream, &ctx->launches);
    CK(cudaGetLastError());
    stage_end(ctx, PGT_STAGE_THERMO);
    pgt_launch_publish(ctx->thermo_out.p, 8, nullptr, 0, ctx->d_small, ctx->stream, &ctx->launches);
    CK(cudaStreamSynchronize(ctx->stream));
    for (int k = 0; k < 4; ++k) out[k] = ctx->h_small[k];
    return PGT_OK;
}

// ---------------------------------------------------------------------------------- exports
static int export_pairs(pgt_ctx* ctx, bool bonds, int64_t cap, int64_t* n, int32_t* oi, int32_t* oj, double* rij_xy, double* fdivr)
{
    if (ctx) RESOLVE();
    if (!ctx || !n) return fail(ctx, PGT_ERR_ARG, "export: null argument");
    if (!ctx->cells_built) return fail(ctx, PGT_ERR_STATE, "export: cells not built");
    CK(cudaSetDevice(ctx->device));
    const int np = ctx->n_live;
    std::vector<int> off((size_t)np + 1), js, orig((size_t)std::max(np, 1));
    std::vector<double2> rb, rcf;
    long long total = 0;
    if (bonds) {
        if (!ctx->bonds_built) return fail(ctx, PGT_ERR_STATE, "export: bonds not built");
        total = ctx->n_hb;
        CK(cudaMemcpy(off.data(), ctx->hb_cnt.p, sizeof(int) * ((size_t)np + 1), cudaMemcpyDeviceToHost));
        js.resize((size_t)total);
        if (total > 0 && cap >= total) {
            rb.resize((size_t)total);
            rcf.resize((size_t)total);
            CK(cudaStreamSynchronize(ctx->stream));
            CK(cudaMemcpy(rb.data(), ctx->hb_b.p, sizeof(double2) * (size_t)total, cudaMemcpyDeviceToHost));
            CK(cudaMemcpy(rcf.data(), ctx->hb_cf.p, sizeof(double2) * (size_t)total, cudaMemcpyDeviceToHost));
            CK(cudaMemcpy(js.data(), ctx->hb_j.p, sizeof(int) * (size_t)total, cudaMemcpyDeviceToHost));
        }
    } else {
        // neighbours within max_rcut: a private count/scan/fill pass (export only)
        DevBuf cnt, jbuf;
        ENSURE(cnt, sizeof(int) * ((size_t)np + 1));
        CK(cudaMemsetAsync(cnt.as<int>() + np, 0, sizeof(int), ctx->stream));
        PgtFrameView fr = frame_view(ctx);
        pgt_launch_bonds(2, ctx->pot, fr, ctx->box, ctx->cells, ctx->max_rcut, cnt.as<int>(), PgtBondOut{}, 1.0, nullptr, ctx->stream,
                         &ctx->launches);
        pgt_launch_scan_exclusive(cnt.as<int>(), cnt.as<int>(), (long long)np + 1, ctx->scan_scratch.as<int>(), ctx->stream, &ctx->launches);
        CK(cudaStreamSynchronize(ctx->stream));
        CK(cudaMemcpy(off.data(), cnt.p, sizeof(int) * ((size_t)np + 1), cudaMemcpyDeviceToHost));
        total = off[np];
        if (total > 0 && cap >= total) {
            ENSURE(jbuf, sizeof(int) * (size_t)total);
            pgt_launch_bonds(3, ctx->pot, fr, ctx->box, ctx->cells, ctx->max_rcut, cnt.as<int>(), PgtBondOut{}, 1.0, jbuf.as<int>(),
                             ctx->stream, &ctx->launches);
            CK(cudaStreamSynchronize(ctx->stream));
            js.resize((size_t)total);
            CK(cudaMemcpy(js.data(), jbuf.p, sizeof(int) * (size_t)total, cudaMemcpyDeviceToHost));
        }
        if (cnt.p) cudaFree(cnt.p);
        if (jbuf.p) cudaFree(jbuf.p);
    }
    *n = total;
    if (cap < total || total == 0) return PGT_OK;
    if (!oi || !oj) return fail(ctx, PGT_ERR_ARG, "export: null output");
    CK(cudaMemcpy(orig.data(), ctx->sorted_orig.p, sizeof(int) * (size_t)np, cudaMemcpyDeviceToHost));
    for (int p = 0; p < np; ++p)
        for (int e = off[p]; e < off[p + 1]; ++e) {
            oi[e] = orig[p];
            oj[e] = orig[js[e]];
            if (bonds && rij_xy) {   // rij = -dx (IrvingKirkwood.h:101)
                rij_xy[2 * e] = -rb[(size_t)e].x;
                rij_xy[2 * e + 1] = -rb[(size_t)e].y;
            }
            if (bonds && fdivr) fdivr[e] = rcf[(size_t)e].y;
        }
    return PGT_OK;
}

int pgt_export_bonds(pgt_ctx* ctx, int64_t cap, int64_t* n, int32_t* i, int32_t* j, double* rij_xy, double* fdivr)
{
    if (ctx && ctx->dim != 2) return fail(ctx, PGT_ERR_UNSUPPORTED, "export: 2-D frames only");
    return export_pairs(ctx, true, cap, n, i, j, rij_xy, fdivr);
}

int pgt_export_neighbors(pgt_ctx* ctx, int64_t cap, int64_t* n, int32_t* i, int32_t* j)
{
    if (ctx && !ctx->have_pot) return fail(ctx, PGT_ERR_STATE, "export: no pair potential");
    if (ctx && ctx->dim != 2) return fail(ctx, PGT_ERR_UNSUPPORTED, "export: 2-D frames only");
    return export_pairs(ctx, false, cap, n, i, j, nullptr, nullptr);
}

int pgt_export_wrapped(pgt_ctx* ctx, double* xy, uint8_t* alive)
{
    if (!ctx || !xy) return fail(ctx, PGT_ERR_ARG, "pgt_export_wrapped: null argument");
    if (!ctx->cells_built) return fail(ctx, PGT_ERR_STATE, "pgt_export_wrapped: cells not built");
    if (ctx->dim != 2) return fail(ctx, PGT_ERR_UNSUPPORTED, "pgt_export_wrapped: 2-D frames only");
    CK(cudaSetDevice(ctx->device));
    RESOLVE();
    const int np = ctx->n_live;
    std::vector<int> orig((size_t)std::max(np, 1));
    std::vector<double> x((size_t)std::max(np, 1)), y((size_t)std::max(np, 1));
    CK(cudaStreamSynchronize(ctx->stream));
    CK(cudaMemcpy(orig.data(), ctx->sorted_orig.p, sizeof(int) * (size_t)np, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(x.data(), ctx->sx.p, sizeof(double) * (size_t)np, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(y.data(), ctx->sy.p, sizeof(double) * (size_t)np, cudaMemcpyDeviceToHost));
    for (long long i = 0; i < ctx->N; ++i) {
        xy[2 * i] = xy[2 * i + 1] = NAN;
        if (alive) alive[i] = 0;
    }
    for (int p = 0; p < np; ++p) {
        xy[2 * (size_t)orig[p]] = x[p];
        xy[2 * (size_t)orig[p] + 1] = y[p];
        if (alive) alive[orig[p]] = 1;
    }
    return PGT_OK;
}

int pgt_stage_time_ms(pgt_ctx* ctx, int stage, float* ms)
{
    if (!ctx || !ms || stage < 0 || stage >= PGT_STAGE_COUNT) return fail(ctx, PGT_ERR_ARG, "pgt_stage_time_ms: bad argument");
    if (!ctx->ev_valid[stage]) return fail(ctx, PGT_ERR_STATE, "stage %d has not run", stage);
    CK(cudaEventSynchronize(ctx->ev1[stage]));
    CK(cudaEventElapsedTime(ms, ctx->ev0[stage], ctx->ev1[stage]));
    return PGT_OK;
}

// scalar host getters ---------------------------------------------------------------------
double pgt_pair_eval(int model_id, double scaled_rcut, const double* params, int nparams, const double rij[3], double di,
                     double dj, int what)
{
    PgtPot P{};
    if (fill_pot(P, model_id, scaled_rcut, params, nparams) != 0) return NAN;
    const double rsq = rij[0] * rij[0] + rij[1] * rij[1] + rij[2] * rij[2];
    double r = NAN;
    PGT_DISPATCH_MODEL(model_id, (r = pgt_pair_eval_t<M>(P, rsq, di, dj, what)));
    return r;
}

double pgt_cg_delta(int kernel_id, double cg_rcut, double dr_sq)
{
    PgtCG cg{};
    cg.kernel = kernel_id;
    cg.rcut = cg_rcut;
    cg.rcutsq = cg_rcut * cg_rcut;
    cg.norm = pgt_cg_norm(kernel_id, cg_rcut);
    return pgt_cg_delta_exact(cg, dr_sq);
}

double pgt_cg_bond(int kernel_id, double cg_rcut, int order, const double x[3], const double ri[3], const double rij[3])
{
    if (order < 1 || order > PGT_MAX_GL) return NAN;
    std::vector<double> gx, gw;
    gauss_legendre_half(order, gx, gw);
    const int m = (order + 1) / 2;
    auto f = [&](double s) {
        double d2 = 0;
        for (int k = 0; k < 3; ++k) {
            const double d = x[k] - (ri[k] + s * rij[k]);
            d2 += d * d;
        }
        return pgt_cg_delta(kernel_id, cg_rcut, d2);
    };
    double sum = 0.0;
    int i0 = 0;
    if (order & 1) { sum = gw[0] * f(0.5); i0 = 1; }
    for (int i = i0; i < m; ++i) {
        const double Ax = 0.5 * gx[i];
        sum += gw[i] * (f(0.5 + Ax) + f(0.5 - Ax));
    }
    return 0.5 * sum;
}

double pgt_cg_bond_adaptive(int kernel_id, double cg_rcut, double epsabs, double epsrel, int limit, const double x[3],
                            const double ri[3], const double rij[3], int* status)
{
    struct Integrand {                                     // getObjFunc (CoarseGrainFunction.h:42-45)
        int kernel_id; double cg_rcut; const double *x, *ri, *rij;
        double operator()(double s) const
        {
            double d2 = 0;
            for (int k = 0; k < 3; ++k) {
                const double d = x[k] - (ri[k] + s * rij[k]);
                d2 += d * d;
            }
            return pgt_cg_delta(kernel_id, cg_rcut, d2);
        }
    };
    int st = 0;
    const double v = pgt_qag15(Integrand{ kernel_id, cg_rcut, x, ri, rij }, 0.0, 1.0, epsabs, epsrel, limit, &st);
    if (status) *status = st;
    return v;
}

int pgt_fp64_peak_probe(pgt_ctx* ctx, double* tflops)
{
    if (!ctx || !tflops) return PGT_ERR_ARG;
    CK(cudaSetDevice(ctx->device));
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, ctx->device));
    const int blocks = prop.multiProcessorCount * 8, threads = 256, iters = 20000;
    DevBuf out;
    ENSURE(out, sizeof(double) * (size_t)blocks * threads);
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    pgt_launch_fp64_probe(out.as<double>(), 1000, blocks, threads, ctx->stream, &ctx->launches);   // warm-up
    float best = 1e30f;
    for (int rep = 0; rep < 3; ++rep) {
        cudaEventRecord(a, ctx->stream);
        pgt_launch_fp64_probe(out.as<double>(), iters, blocks, threads, ctx->stream, &ctx->launches);
        cudaEventRecord(b, ctx->stream);
        CK(cudaEventSynchronize(b));
        float ms = 0;
        cudaEventElapsedTime(&ms, a, b);
        best = std::min(best, ms);
    }
    cudaEventDestroy(a);
    cudaEventDestroy(b);
    cudaFree(out.p);
    *tflops = 2.0 * 8.0 * (double)iters * blocks * threads / (best * 1e-3) / 1e12;
    return PGT_OK;
}

// ---------------------------------------------------------------------------------- NCCL
int pgt_comm_unique_id(void* id128)
{
    std::string err;
    if (!id128) return PGT_ERR_ARG;
    if (!load_nccl(err)) { g_create_error = err; return PGT_ERR_NCCL; }
    pgtNcclUniqueId id;
    int rc = g_nccl.GetUniqueId(&id);
    if (rc != 0) { g_create_error = g_nccl.GetErrorString(rc); return PGT_ERR_NCCL; }
    memcpy(id128, &id, 128);
    return PGT_OK;
}

int pgt_comm_init(pgt_ctx* ctx, int nranks, int rank, const void* id128)
{
    if (!ctx || !id128 || nranks < 1 || rank < 0 || rank >= nranks) return fail(ctx, PGT_ERR_ARG, "pgt_comm_init: bad argument");
    std::string err;
    if (!load_nccl(err)) return fail(ctx, PGT_ERR_NCCL, "%s", err.c_str());
    CK(cudaSetDevice(ctx->device));
    pgtNcclUniqueId id;
    memcpy(&id, id128, 128);
    int rc = g_nccl.CommInitRank(&ctx->comm, nranks, id, rank);
    if (rc != 0) return fail(ctx, PGT_ERR_NCCL, "ncclCommInitRank: %s", g_nccl.GetErrorString(rc));
    ctx->nranks = nranks;
    ctx->rank = rank;
    return PGT_OK;
}

int pgt_comm_allgather_field(pgt_ctx* ctx, int field_id, const int64_t* counts, void* out_dev)
{
    if (!ctx || !counts || !out_dev) return fail(ctx, PGT_ERR_ARG, "pgt_comm_allgather_field: null argument");
    if (!ctx->comm) return fail(ctx, PGT_ERR_STATE, "pgt_comm_allgather_field: communicator not initialised");
    DevBuf* b;
    int nc;
    int rc = field_info(ctx, field_id, &b, &nc);
    if (rc != 0) return fail(ctx, PGT_ERR_STATE, "field %d not available", field_id);
    if (counts[ctx->rank] != (int64_t)nc * (ctx->field_g1 - ctx->field_g0))
        return fail(ctx, PGT_ERR_ARG, "pgt_comm_allgather_field: counts[%d] = %lld, the local field holds %lld doubles", ctx->rank,
                    (long long)counts[ctx->rank], (long long)nc * (ctx->field_g1 - ctx->field_g0));
    CK(cudaSetDevice(ctx->device));
    if (ctx->dl_pending) CK(cudaStreamWaitEvent(ctx->stream, ctx->ev_dl_done, 0));   // a read-back of this field may still be in flight
    // variable-size all-gather = one broadcast per rank inside a group (NVLink / NVSwitch transport)
    const int ncclDouble = 8;
    int nrc = g_nccl.GroupStart();
    size_t off = 0;
    for (int r = 0; r < ctx->nranks && nrc == 0; ++r) {
        double* dst = reinterpret_cast<double*>(out_dev) + off;
        const void* src = (r == ctx->rank) ? b->p : (const void*)dst;
        if (counts[r] > 0) nrc = g_nccl.Broadcast(src, dst, (size_t)counts[r], ncclDouble, r, ctx->comm, ctx->stream);
        off += (size_t)counts[r];
    }
    int nrc2 = g_nccl.GroupEnd();
    if (nrc != 0 || nrc2 != 0) return fail(ctx, PGT_ERR_NCCL, "ncclBroadcast group: %s", g_nccl.GetErrorString(nrc ? nrc : nrc2));
    return PGT_OK;
}

int pgt_comm_allgather_frame(pgt_ctx* ctx)
{
    if (!ctx) return PGT_ERR_ARG;
    if (!ctx->comm) return fail(ctx, PGT_ERR_STATE, "pgt_comm_allgather_frame: communicator not initialised");
    if (!ctx->have_frame || ctx->part_chunk <= 0) return fail(ctx, PGT_ERR_STATE, "pgt_comm_allgather_frame: no frame part uploaded (pgt_frame_upload_part)");
    CK(cudaSetDevice(ctx->device));
    const size_t es = ctx->is_f32 ? 4 : 8;
    const int dtype = ctx->is_f32 ? 7 /* ncclFloat32 */ : 8 /* ncclFloat64 */;
    const size_t chunk = (size_t)ctx->part_chunk;
    if (ctx->raw_cur < 0) return fail(ctx, PGT_ERR_STATE, "pgt_comm_allgather_frame: the frame was not uploaded by pgt_frame_upload_part");
    wait_upload(ctx);
    char* pos = static_cast<char*>(ctx->raw_pos2[ctx->raw_cur].p);
    char* diam = static_cast<char*>(ctx->raw_diam2[ctx->raw_cur].p);
    // in place: this rank's part already sits at its offset of the receive buffer
    int nrc = g_nccl.GroupStart();
    if (nrc == 0) nrc = g_nccl.AllGather(pos + (size_t)ctx->rank * chunk * 3 * es, pos, chunk * 3, dtype, ctx->comm, ctx->stream);
    if (nrc == 0) nrc = g_nccl.AllGather(diam + (size_t)ctx->rank * chunk * es, diam, chunk, dtype, ctx->comm, ctx->stream);
    const int nrc2 = g_nccl.GroupEnd();
    if (nrc != 0 || nrc2 != 0) return fail(ctx, PGT_ERR_NCCL, "ncclAllGather: %s", g_nccl.GetErrorString(nrc ? nrc : nrc2));
    return PGT_OK;
}

int pgt_thermo_set_shard(pgt_ctx* ctx, int part, int nparts)
{
    if (!ctx) return PGT_ERR_ARG;
    if (nparts < 1 || part < 0 || part >= nparts) return fail(ctx, PGT_ERR_ARG, "pgt_thermo_set_shard: part %d of %d", part, nparts);
    ctx->thermo_part = part;
    ctx->thermo_nparts = nparts;
    return PGT_OK;
}

int pgt_comm_allreduce_sum(pgt_ctx* ctx, double* inout_host, int n)
{
    if (!ctx || !inout_host || n < 0 || n > 256) return fail(ctx, PGT_ERR_ARG, "pgt_comm_allreduce_sum: bad argument");
    if (!ctx->comm) return fail(ctx, PGT_ERR_STATE, "pgt_comm_allreduce_sum: communicator not initialised");
    CK(cudaSetDevice(ctx->device));
    ENSURE(ctx->comm_scratch, sizeof(double) * 256);
    memcpy(ctx->h_small, inout_host, sizeof(double) * n);
    CK(cudaMemcpyAsync(ctx->comm_scratch.p, ctx->h_small, sizeof(double) * n, cudaMemcpyHostToDevice, ctx->stream));
    const int ncclDouble = 8, ncclSum = 0;
    int rc = g_nccl.AllReduce(ctx->comm_scratch.p, ctx->comm_scratch.p, (size_t)n, ncclDouble, ncclSum, ctx->comm, ctx->stream);
    if (rc != 0) return fail(ctx, PGT_ERR_NCCL, "ncclAllReduce: %s", g_nccl.GetErrorString(rc));
    CK(cudaMemcpyAsync(ctx->h_small, ctx->comm_scratch.p, sizeof(double) * n, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    memcpy(inout_host, ctx->h_small, sizeof(double) * n);
    return PGT_OK;
}

int pgt_comm_destroy(pgt_ctx* ctx)
{
    if (!ctx) return PGT_ERR_ARG;
    if (ctx->comm && g_nccl.ok) g_nccl.CommDestroy(ctx->comm);
    ctx->comm = nullptr;
    ctx->nranks = 1;
    ctx->rank = 0;
    return PGT_OK;
}

}  // extern "C"
