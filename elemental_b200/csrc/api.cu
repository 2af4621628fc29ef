// extern "C" surface of libhydrogen_b200.so (declared in include/hydrogen_b200.h).
#include "../../include/hydrogen_b200.h"
#include "grid.h"
#include "hyd_internal.h"
#include <string.h>
#include <stdio.h>
#include <unistd.h>
#include <sys/stat.h>
#include <time.h>
#include <stdlib.h>
#include <algorithm>

namespace hyd {
const char* last_error();
int probe_peak(int kind, int iters, int ctas_per_sm, double* tflops);
template <typename T>
int dist_gemm(Grid& g, int oA, int oB, int alg, T alpha, const DM& A, const DM& B, T beta, const DM& C, i64 blocksize,
              cudaStream_t s, HostImages* host);
int dist_gemm_half(Grid& g, int bf16, int oA, int oB, int alg, float alpha, const DM& A, const DM& B, float beta, const DM& C,
                   i64 blocksize, cudaStream_t s);
struct h16;
template <typename T>
int redistribute(Grid& g, const DM& src, const DM& dst, bool transpose, int slot, cudaStream_t s);
template <typename T>
int dist_trrk(Grid& g, int uplo, int oA, int oB, T alpha, const DM& A, const DM& B, T beta, const DM& C, i64 blocksize,
              cudaStream_t s);
template <typename T>
int dist_make_symmetric(Grid& g, int uplo, const DM& A, cudaStream_t s);
template <typename T>
int dist_symm(Grid& g, int side, int uplo, T alpha, const DM& A, const DM& B, T beta, const DM& C, i64 blocksize, cudaStream_t s);
template <typename T>
int dist_trapezoid(Grid& g, int op, int uplo, T alpha, const DM* X, const DM& Y, i64 offset, cudaStream_t s);
}  // namespace hyd

struct hyd_world { hyd::World* w; };
struct hyd_grid { hyd::Grid* g; };
struct hyd_local_group { hyd::LocalShared* sh; };

using hyd::set_error;

static hyd::DM to_dm(const hyd_dm* a) {
    hyd::DM d;
    d.buf = a->buf; d.ldim = a->ldim; d.height = a->height; d.width = a->width;
    d.colDist = a->colDist; d.rowDist = a->rowDist; d.colAlign = a->colAlign; d.rowAlign = a->rowAlign; d.root = a->root;
    return d;
}
static void from_dm(const hyd::DM& d, hyd_dm* a) {
    a->buf = d.buf; a->ldim = d.ldim; a->height = d.height; a->width = d.width;
    a->colDist = d.colDist; a->rowDist = d.rowDist; a->colAlign = d.colAlign; a->rowAlign = d.rowAlign; a->root = d.root;
    a->reserved = 0;
}
static int check_dm(const hyd::GridShape& sh, const hyd_dm* a, const char* who) {
    if (!a) return set_error("%s: null matrix", who);
    if (a->height < 0 || a->width < 0) return set_error("%s: negative dimension", who);
    if (!hyd::pair_supported(a->colDist, a->rowDist))
        return set_error("%s: unsupported distribution pair [%d,%d]", who, a->colDist, a->rowDist);
    if (a->colAlign < 0 || a->colAlign >= sh.stride(a->colDist) || a->rowAlign < 0 || a->rowAlign >= sh.stride(a->rowDist))
        return set_error("%s: alignment out of range", who);
    return 0;
}

// ---- host-buffer entry points (the call a CPU-resident caller makes) ----
template <typename T>
static int gemm_host(hyd_grid* g, int oA, int oB, int alg, T alpha, const hyd_dm* A, const hyd_dm* B, T beta,
                     const hyd_dm* C, int64_t blocksize, cudaStream_t s) {
    hyd::Grid& G = *g->g;
    const hyd::GridShape& sh = G.shape;
    if (check_dm(sh, A, "hyd_gemm_host(A)") || check_dm(sh, B, "hyd_gemm_host(B)") || check_dm(sh, C, "hyd_gemm_host(C)"))
        return 1;
    const hyd_dm* host[3] = {A, B, C};
    hyd::DM dev[3];
    hyd::LocalShape L[3];
    for (int i = 0; i < 3; ++i) {
        dev[i] = to_dm(host[i]);
        L[i] = hyd::local_shape(sh, dev[i], G.mc, G.mr);
        const int64_t q = 16 / (int64_t)sizeof(T);
        dev[i].ldim = L[i].lh > 0 ? (L[i].lh + q - 1) / q * q : q;
        const size_t bytes = (size_t)dev[i].ldim * (size_t)(L[i].lw > 0 ? L[i].lw : 1) * sizeof(T);
        if (hyd::ws_reserve(G.host_stage[i], bytes)) return 1;
        dev[i].buf = G.host_stage[i].buf;
        if (L[i].lh > 0 && L[i].lw > 0) {
            if (!host[i]->buf || host[i]->ldim < L[i].lh) return set_error("hyd_gemm_host: bad host buffer / ldim");
            if (i == 2 && beta == T(0)) continue;  // C is overwritten: no upload
            if (i < 2) continue;                   // A, B: uploaded by dist_gemm (panel-pipelined for SUMMA-C)
            if (hyd::check_cuda(cudaMemcpy2DAsync(dev[i].buf, (size_t)dev[i].ldim * sizeof(T), host[i]->buf,
                                                  (size_t)host[i]->ldim * sizeof(T), (size_t)L[i].lh * sizeof(T),
                                                  (size_t)L[i].lw, cudaMemcpyHostToDevice, s),
                                "hyd_gemm_host: H2D"))
                return 1;
        }
    }
    // c_downloaded: set when the engine already wrote C's blocks to the host (column-blocked path)
    hyd::HostImages hi{A->buf, A->ldim, B->buf, B->ldim, C->buf, C->ldim, 0};
    if (hyd::dist_gemm<T>(G, oA, oB, alg, alpha, dev[0], dev[1], beta, dev[2], blocksize, s, &hi)) return 1;
    if (!hi.c_downloaded && L[2].lh > 0 && L[2].lw > 0 &&
        hyd::check_cuda(cudaMemcpy2DAsync(C->buf, (size_t)C->ldim * sizeof(T), dev[2].buf, (size_t)dev[2].ldim * sizeof(T),
                                          (size_t)L[2].lh * sizeof(T), (size_t)L[2].lw, cudaMemcpyDeviceToHost, s),
                        "hyd_gemm_host: D2H"))
        return 1;
    return hyd::check_cuda(hyd::stream_sync(s), "hyd_gemm_host: sync");
}
extern "C" {

const char* hyd_last_error(void) { return hyd::last_error(); }
int hyd_version(void) { return 100; }
long long hyd_launch_count(void) { return hyd::g_launches.load(); }
long long hyd_alloc_count(void) { return hyd::g_allocs.load(); }
long long hyd_sync_count(void) { return hyd::g_syncs.load(); }

int64_t hyd_mod(int64_t a, int64_t b) { return hyd::mod_(a, b); }
int64_t hyd_shift(int64_t rank, int64_t align, int64_t stride) { return hyd::shift_(rank, align, stride); }
int64_t hyd_length(int64_t n, int64_t shift, int64_t stride) { return hyd::length_(n, shift, stride); }
int64_t hyd_max_length(int64_t n, int64_t stride) { return hyd::max_length_(n, stride); }

// ---- local kernels ----
int hyd_lgemm_f64(int tA, int tB, int64_t m, int64_t n, int64_t k, double alpha, const double* A, int64_t lda,
                  const double* B, int64_t ldb, double beta, double* C, int64_t ldc, hyd_stream_t stream) {
    ++hyd::g_launches;
    return hyd::lgemm_f64(tA ? 1 : 0, tB ? 1 : 0, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc, (cudaStream_t)stream);
}
int hyd_lgemm_f32(int tA, int tB, int64_t m, int64_t n, int64_t k, float alpha, const float* A, int64_t lda,
                  const float* B, int64_t ldb, float beta, float* C, int64_t ldc, hyd_stream_t stream) {
    ++hyd::g_launches;
    return hyd::lgemm_f32(tA ? 1 : 0, tB ? 1 : 0, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc, (cudaStream_t)stream);
}
int hyd_lgemm_f16(int tA, int tB, int64_t m, int64_t n, int64_t k, float alpha, const void* A, int64_t lda, const void* B,
                  int64_t ldb, float beta, void* C, int64_t ldc, hyd_stream_t stream) {
    ++hyd::g_launches;
    return hyd::lgemm_f16(tA ? 1 : 0, tB ? 1 : 0, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc, (cudaStream_t)stream);
}
int hyd_lgemm_bf16(int tA, int tB, int64_t m, int64_t n, int64_t k, float alpha, const void* A, int64_t lda, const void* B,
                   int64_t ldb, float beta, void* C, int64_t ldc, hyd_stream_t stream) {
    ++hyd::g_launches;
    return hyd::lgemm_bf16(tA ? 1 : 0, tB ? 1 : 0, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc, (cudaStream_t)stream);
}
int hyd_probe_peak(int kind, int iters, int ctas_per_sm, double* tflops) {
    return hyd::probe_peak(kind, iters, ctas_per_sm, tflops);
}

// ---- world / grid ----
int hyd_nccl_unique_id(void* out128) { return hyd::nccl_unique_id(out128); }
int hyd_world_create_nccl(const void* id128, int rank, int nranks, hyd_world** out) {
    if (!id128 || !out || nranks <= 0 || rank < 0 || rank >= nranks) return set_error("hyd_world_create_nccl: bad arguments");
    hyd::World* w = hyd::make_nccl_world(id128, rank, nranks);
    if (!w) return 1;
    *out = new hyd_world{w};
    return 0;
}
// One process per GPU without any launcher library: rank / size / local rank from the environment the usual launchers export,
// the NCCL id through a rendezvous file (one node: the scope of this library's transports).
static int env_int(const char* const* names, int fallback) {
    for (int i = 0; names[i]; ++i)
        if (const char* e = getenv(names[i])) return atoi(e);
    return fallback;
}
// host-only pieces of hyd_world_create_env (exported: tests/test_multirank_cpu.py runs them in two processes without a GPU)
int hyd_env_world_info(int* rank, int* size, int* local_rank) {
    static const char* const kRank[] = {"RANK", "OMPI_COMM_WORLD_RANK", "PMI_RANK", "SLURM_PROCID", nullptr};
    static const char* const kSize[] = {"WORLD_SIZE", "OMPI_COMM_WORLD_SIZE", "PMI_SIZE", "SLURM_NTASKS", nullptr};
    static const char* const kLocal[] = {"LOCAL_RANK", "OMPI_COMM_WORLD_LOCAL_RANK", "MPI_LOCALRANKID", "SLURM_LOCALID", nullptr};
    const int r = env_int(kRank, 0), n = env_int(kSize, 1);
    if (n <= 0 || r < 0 || r >= n) return set_error("hyd_world_create_env: rank %d of %d from the environment", r, n);
    if (rank) *rank = r;
    if (size) *size = n;
    if (local_rank) *local_rank = env_int(kLocal, r);
    return 0;
}
// Rendezvous file, one per launch on this node.  The launch is identified by what its ranks share: the launcher's job id when
// there is one (torchrun's TORCHELASTIC_RUN_ID, SLURM_JOB_ID, PMIX_NAMESPACE) plus MASTER_PORT, else the SESSION id -- not the
// parent pid: wrappers such as `timeout`, `env` or `numactl` give every rank a parent of its own (found the hard way: two ranks
// under `timeout` waited for two different files), while the session survives them.  HYD_RENDEZVOUS_FILE overrides.
static void env_rendezvous_path(char* path, size_t cap) {
    if (const char* f = getenv("HYD_RENDEZVOUS_FILE")) { snprintf(path, cap, "%s", f); return; }
    const char* dir = getenv("HYD_RENDEZVOUS_DIR");
    const char* port = getenv("MASTER_PORT");
    const char* job = getenv("TORCHELASTIC_RUN_ID");
    if (!job || !*job || !strcmp(job, "none")) job = getenv("SLURM_JOB_ID");
    if (!job || !*job) job = getenv("PMIX_NAMESPACE");
    char key[160];
    if (job && *job) {
        snprintf(key, sizeof(key), "job%.120s", job);
        for (char* c = key; *c; ++c) if (*c == '/' || *c == ' ') *c = '_';
    } else {
        snprintf(key, sizeof(key), "sid%d", (int)getsid(0));
    }
    snprintf(path, cap, "%s/hyd_rdv_%d_%s_%s", dir ? dir : "/tmp", (int)getuid(), port ? port : "0", key);
}
// rank 0 publishes its 128 bytes (written to a temporary name, then renamed: readers never see a partial file); every other
// rank polls for the file (up to timeout_s seconds) and reads them
int hyd_env_exchange_id(int rank, void* id128, int timeout_s) {
    if (!id128 || rank < 0) return set_error("hyd_env_exchange_id: bad arguments");
    char path[512];
    env_rendezvous_path(path, sizeof(path));
    if (rank == 0) {
        char tmp[600];
        snprintf(tmp, sizeof(tmp), "%s.tmp%d", path, (int)getpid());
        FILE* f = fopen(tmp, "wb");
        if (!f || fwrite(id128, 1, 128, f) != 128) { if (f) fclose(f); return set_error("hyd_world_create_env: cannot write %s", tmp); }
        fclose(f);
        if (rename(tmp, path) != 0) return set_error("hyd_world_create_env: cannot publish %s", path);
        return 0;
    }
    for (int tries = 0; tries < 10 * (timeout_s > 0 ? timeout_s : 120); ++tries) {
        struct stat st;
        // a file older than a minute is the leftover of a launch that died between publishing and joining: not ours
        if (stat(path, &st) == 0 && time(nullptr) - st.st_mtime <= 60) {
            FILE* f = fopen(path, "rb");
            if (f) {
                const bool got = fread(id128, 1, 128, f) == 128;
                fclose(f);
                if (got) return 0;
            }
        }
        usleep(100000);
    }
    return set_error("hyd_world_create_env: rank 0 never published %s", path);
}
// rank 0, once every rank has read the id: remove the file
int hyd_env_exchange_done(int rank) {
    if (rank == 0) {
        char path[512];
        env_rendezvous_path(path, sizeof(path));
        unlink(path);
    }
    return 0;
}
int hyd_world_create_env(hyd_world** out) {
    if (!out) return set_error("hyd_world_create_env: bad arguments");
    int rank = 0, size = 1, local = 0;
    if (hyd_env_world_info(&rank, &size, &local)) return 1;
    int ndev = 0;
    if (hyd::check_cuda(cudaGetDeviceCount(&ndev), "hyd_world_create_env") || ndev <= 0) return set_error("hyd_world_create_env: no CUDA device");
    if (hyd::check_cuda(cudaSetDevice(local % ndev), "hyd_world_create_env: cudaSetDevice")) return 1;
    unsigned char id[128];
    if (rank == 0 && hyd::nccl_unique_id(id)) return 1;
    if (hyd_env_exchange_id(rank, id, 120)) return 1;
    hyd::World* w = hyd::make_nccl_world(id, rank, size);
    if (!w) { hyd_env_exchange_done(rank); return 1; }
    const int rc = w->barrier();   // every rank has read the id
    hyd_env_exchange_done(rank);
    if (rc) { delete w; return 1; }
    *out = new hyd_world{w};
    return 0;
}
const char* hyd_world_kind(const hyd_world* w) { return w && w->w ? w->w->kind() : "none"; }
int hyd_world_size(const hyd_world* w) { return w && w->w ? w->w->size : 0; }
int hyd_world_rank(const hyd_world* w) { return w && w->w ? w->w->rank : -1; }
int hyd_grid_default_height(int size) {
    if (size <= 0) return 0;
    int h = 1;
    while ((long long)(h + 1) * (h + 1) <= size) ++h;   // floor(sqrt(size)) without floating point
    while (size % h != 0) ++h;
    return h;
}
int hyd_local_group_create(int nranks, hyd_local_group** out) {
    if (nranks <= 0 || !out) return set_error("hyd_local_group_create: bad arguments");
    *out = new hyd_local_group{hyd::local_shared_create(nranks)};
    return 0;
}
int hyd_local_group_destroy(hyd_local_group* grp) {
    if (grp) { hyd::local_shared_destroy(grp->sh); delete grp; }
    return 0;
}
int hyd_world_create_local(hyd_local_group* grp, int rank, hyd_world** out) {
    if (!grp || !out) return set_error("hyd_world_create_local: bad arguments");
    *out = new hyd_world{hyd::make_local_world(grp->sh, rank)};
    return 0;
}
int hyd_world_destroy(hyd_world* w) {
    if (w) { delete w->w; delete w; }
    return 0;
}
int hyd_world_barrier(hyd_world* w) { return w ? w->w->barrier() : set_error("hyd_world_barrier: null world"); }

int hyd_grid_create(hyd_world* w, int height, int order, hyd_grid** out) {
    if (!w || !out) return set_error("hyd_grid_create: bad arguments");
    const int size = w->w->size;
    if (height <= 0) height = hyd_grid_default_height(size);   // Grid.cpp:87: height <= 0 ? DefaultHeight(size) : height
    if (height <= 0 || size % height != 0)
        return set_error("Grid height, %d, does not evenly divide grid size, %d", height, size);  // Grid.cpp:98-100
    hyd::Grid* g = new hyd::Grid;
    g->world = w->w;
    g->shape.r = height;
    g->shape.c = size / height;
    g->shape.col_major = order ? 1 : 0;
    g->rank = w->w->rank;
    g->shape.coords(g->rank, g->mc, g->mr);
    if (const char* e = getenv("HYD_HOST_CBLOCKS")) g->host_cblocks = std::max(1, std::min(8, atoi(e)));
    *out = new hyd_grid{g};
    return 0;
}
int hyd_grid_destroy(hyd_grid* g) {
    if (g) { hyd::grid_destroy(g->g); delete g; }
    return 0;
}
int hyd_grid_height(const hyd_grid* g) { return g->g->shape.r; }
int hyd_grid_width(const hyd_grid* g) { return g->g->shape.c; }
int hyd_grid_size(const hyd_grid* g) { return g->g->shape.size(); }
int hyd_grid_rank(const hyd_grid* g) { return g->g->rank; }
int hyd_grid_row(const hyd_grid* g) { return g->g->mc; }
int hyd_grid_col(const hyd_grid* g) { return g->g->mr; }
int hyd_grid_vc_rank(const hyd_grid* g) { return g->g->vc(); }
int hyd_grid_vr_rank(const hyd_grid* g) { return g->g->vr(); }
int hyd_grid_set_overlap(hyd_grid* g, int overlap, int sm_reserve) {
    g->g->overlap = overlap;
    if (sm_reserve >= 0) g->g->sm_reserve = sm_reserve;
    return 0;
}
int hyd_grid_set_host_blocks(hyd_grid* g, int max_blocks) {
    if (!g || max_blocks < 1) return set_error("hyd_grid_set_host_blocks: bad arguments");
    g->g->host_cblocks = max_blocks > 8 ? 8 : max_blocks;
    return 0;
}
int hyd_grid_profile(hyd_grid* g, int enable) {
    hyd::Grid& G = *g->g;
    if (hyd::check_cuda(cudaDeviceSynchronize(), "hyd_grid_profile")) return 1;
    G.profile = enable != 0;
    for (auto& u : G.prof_used) u = 0;
    return 0;
}
int hyd_grid_profile_read(hyd_grid* g, int kind, long long* launches, double* total_ms) {
    hyd::Grid& G = *g->g;
    if (kind < 0 || kind >= hyd::Grid::PROF_KINDS) return set_error("hyd_grid_profile_read: bad kind");
    if (hyd::check_cuda(cudaDeviceSynchronize(), "hyd_grid_profile_read")) return 1;
    double tot = 0;
    for (size_t i = 0; i + 1 < G.prof_used[kind]; i += 2) {
        float ms = 0;
        if (hyd::check_cuda(cudaEventElapsedTime(&ms, G.prof_ev[kind][i], G.prof_ev[kind][i + 1]), "cudaEventElapsedTime")) return 1;
        tot += ms;
    }
    if (launches) *launches = (long long)(G.prof_used[kind] / 2);
    if (total_ms) *total_ms = tot;
    return 0;
}
int hyd_grid_profile_trace(hyd_grid* g, int kind, int ref_kind, double* start_ms, double* end_ms, int max_entries) {
    hyd::Grid& G = *g->g;
    if (kind < 0 || kind >= hyd::Grid::PROF_KINDS || ref_kind < 0 || ref_kind >= hyd::Grid::PROF_KINDS || !start_ms || !end_ms)
        return set_error("hyd_grid_profile_trace: bad arguments"), -1;
    if (hyd::check_cuda(cudaDeviceSynchronize(), "hyd_grid_profile_trace")) return -1;
    if (G.prof_used[ref_kind] < 2) return 0;
    cudaEvent_t t0 = G.prof_ev[ref_kind][0];
    int n = 0;
    for (size_t i = 0; i + 1 < G.prof_used[kind] && n < max_entries; i += 2, ++n) {
        float a = 0, b = 0;
        if (cudaEventElapsedTime(&a, t0, G.prof_ev[kind][i]) != cudaSuccess || cudaEventElapsedTime(&b, t0, G.prof_ev[kind][i + 1]) != cudaSuccess) {
            cudaGetLastError();
            break;
        }
        start_ms[n] = a;
        end_ms[n] = b;
    }
    return n;
}
int hyd_dist_stride(const hyd_grid* g, int dist) { return g->g->shape.stride(dist); }
int hyd_dist_rank(const hyd_grid* g, int dist) { return g->g->shape.dist_rank(dist, g->g->mc, g->g->mr); }

int hyd_local_shape_of(int gh, int gw, int order, int rank, const hyd_dm* a, int64_t* cs, int64_t* rs, int64_t* lh,
                       int64_t* lw) {
    hyd::GridShape sh;
    sh.r = gh; sh.c = gw; sh.col_major = order ? 1 : 0;
    if (gh <= 0 || gw <= 0 || rank < 0 || rank >= gh * gw) return set_error("hyd_local_shape_of: bad grid/rank");
    if (check_dm(sh, a, "hyd_local_shape_of")) return 1;
    int mc, mr;
    sh.coords(rank, mc, mr);
    const hyd::LocalShape L = hyd::local_shape(sh, to_dm(a), mc, mr);
    if (cs) *cs = L.colShift;
    if (rs) *rs = L.rowShift;
    if (lh) *lh = L.lh;
    if (lw) *lw = L.lw;
    return 0;
}
int hyd_local_shape(const hyd_grid* g, const hyd_dm* a, int64_t* cs, int64_t* rs, int64_t* lh, int64_t* lw) {
    const hyd::GridShape& sh = g->g->shape;
    return hyd_local_shape_of(sh.r, sh.c, sh.col_major, g->g->rank, a, cs, rs, lh, lw);
}
int hyd_view(const hyd_grid* g, const hyd_dm* a, int64_t i, int64_t j, int64_t h, int64_t w, int es, hyd_dm* out) {
    if (check_dm(g->g->shape, a, "hyd_view")) return 1;
    if (i < 0 || j < 0 || h < 0 || w < 0 || i + h > a->height || j + w > a->width)
        return set_error("hyd_view: range out of bounds");
    from_dm(hyd::view(g->g->shape, to_dm(a), g->g->mc, g->g->mr, i, j, h, w, es), out);
    return 0;
}

int hyd_gemm_default_alg(int64_t m, int64_t n, int64_t k) { return hyd::gemm_default_alg(m, n, k); }
int hyd_adopted_align(int gh, int gw, int srcDist, int srcAlign, int dstDist) {
    hyd::GridShape sh;
    sh.r = gh; sh.c = gw;
    return hyd::adopted_align(sh, srcDist, srcAlign, dstDist);
}
int hyd_host_block_plan(int64_t tm, int64_t tn, int64_t sms, double lambda, double mu, int max_blocks, int64_t* cuts) {
    if (!cuts || max_blocks < 1) return 0;
    hyd::i64 c[9];
    const int nb = hyd::host_col_cuts(tm, tn, sms, lambda, mu, max_blocks, c);
    for (int i = 0; i <= nb; ++i) cuts[i] = c[i];
    return nb;
}
int hyd_trrk_plan(int gh, int gw, int order, int rank, const hyd_dm* C, int uplo, int64_t j0, int64_t w, int es, int64_t* cuts) {
    hyd::GridShape sh;
    sh.r = gh; sh.c = gw; sh.col_major = order ? 1 : 0;
    if (gh <= 0 || gw <= 0 || rank < 0 || rank >= gh * gw || !cuts) return set_error("hyd_trrk_plan: bad arguments");
    if ((uplo != 0 && uplo != 1) || (es != 4 && es != 8)) return set_error("hyd_trrk_plan: bad uplo / element size");
    if (check_dm(sh, C, "hyd_trrk_plan")) return 1;
    int mc, mr;
    sh.coords(rank, mc, mr);
    const hyd::LocalShape L = hyd::local_shape(sh, to_dm(C), mc, mr);
    if (j0 < 0 || w <= 0 || j0 + w > L.lw) return set_error("hyd_trrk_plan: column block outside the local block");
    const hyd::TrrkCuts k = hyd::trrk_cuts(uplo, L, j0, w, 16 / es);
    cuts[0] = k.d0; cuts[1] = k.d1; cuts[2] = k.t0; cuts[3] = k.t1;
    return 0;
}

int hyd_redist_plan(int gh, int gw, int order, int rank, const hyd_dm* src, const hyd_dm* dst, int transpose, int es,
                    int64_t* sendBytes, int64_t* recvBytes) {
    hyd::GridShape sh;
    sh.r = gh; sh.c = gw; sh.col_major = order ? 1 : 0;
    if (gh <= 0 || gw <= 0 || rank < 0 || rank >= gh * gw || !sendBytes || !recvBytes) return set_error("hyd_redist_plan: bad arguments");
    if (check_dm(sh, src, "hyd_redist_plan") || check_dm(sh, dst, "hyd_redist_plan")) return 1;
    return hyd::redist_plan(sh, rank, to_dm(src), to_dm(dst), transpose != 0, es, (hyd::i64*)sendBytes, (hyd::i64*)recvBytes);
}

// ---- distributed operations ----
int hyd_redist_f64(hyd_grid* g, const hyd_dm* src, const hyd_dm* dst, int transpose, hyd_stream_t stream) {
    if (check_dm(g->g->shape, src, "hyd_redist") || check_dm(g->g->shape, dst, "hyd_redist")) return 1;
    return hyd::redistribute<double>(*g->g, to_dm(src), to_dm(dst), transpose != 0, 3, (cudaStream_t)stream);
}
int hyd_redist_f32(hyd_grid* g, const hyd_dm* src, const hyd_dm* dst, int transpose, hyd_stream_t stream) {
    if (check_dm(g->g->shape, src, "hyd_redist") || check_dm(g->g->shape, dst, "hyd_redist")) return 1;
    return hyd::redistribute<float>(*g->g, to_dm(src), to_dm(dst), transpose != 0, 3, (cudaStream_t)stream);
}
int hyd_axpy_contract_f64(hyd_grid* g, double alpha, const hyd_dm* src, const hyd_dm* dst, int transpose,
                          hyd_stream_t stream) {
    if (check_dm(g->g->shape, src, "hyd_axpy_contract") || check_dm(g->g->shape, dst, "hyd_axpy_contract")) return 1;
    return hyd::reduce_into<double>(*g->g, alpha, to_dm(src), to_dm(dst), transpose != 0, 3, (cudaStream_t)stream);
}
int hyd_axpy_contract_f32(hyd_grid* g, float alpha, const hyd_dm* src, const hyd_dm* dst, int transpose,
                          hyd_stream_t stream) {
    if (check_dm(g->g->shape, src, "hyd_axpy_contract") || check_dm(g->g->shape, dst, "hyd_axpy_contract")) return 1;
    return hyd::reduce_into<float>(*g->g, alpha, to_dm(src), to_dm(dst), transpose != 0, 3, (cudaStream_t)stream);
}
int hyd_gemm_f64(hyd_grid* g, int oA, int oB, int alg, double alpha, const hyd_dm* A, const hyd_dm* B, double beta,
                 const hyd_dm* C, int64_t blocksize, hyd_stream_t stream) {
    const hyd::GridShape& sh = g->g->shape;
    if (check_dm(sh, A, "hyd_gemm(A)") || check_dm(sh, B, "hyd_gemm(B)") || check_dm(sh, C, "hyd_gemm(C)")) return 1;
    return hyd::dist_gemm<double>(*g->g, oA, oB, alg, alpha, to_dm(A), to_dm(B), beta, to_dm(C), blocksize,
                                  (cudaStream_t)stream, nullptr);
}
int hyd_gemm_f32(hyd_grid* g, int oA, int oB, int alg, float alpha, const hyd_dm* A, const hyd_dm* B, float beta,
                 const hyd_dm* C, int64_t blocksize, hyd_stream_t stream) {
    const hyd::GridShape& sh = g->g->shape;
    if (check_dm(sh, A, "hyd_gemm(A)") || check_dm(sh, B, "hyd_gemm(B)") || check_dm(sh, C, "hyd_gemm(C)")) return 1;
    return hyd::dist_gemm<float>(*g->g, oA, oB, alg, alpha, to_dm(A), to_dm(B), beta, to_dm(C), blocksize,
                                 (cudaStream_t)stream, nullptr);
}


// ---- local BLAS-1-like kernels ----
int hyd_transpose_f64(int64_t m, int64_t n, const double* A, int64_t lda, double* B, int64_t ldb, hyd_stream_t s) {
    return hyd::transpose_local<double>(m, n, A, lda, B, ldb, (cudaStream_t)s);
}
int hyd_transpose_f32(int64_t m, int64_t n, const float* A, int64_t lda, float* B, int64_t ldb, hyd_stream_t s) {
    return hyd::transpose_local<float>(m, n, A, lda, B, ldb, (cudaStream_t)s);
}
int hyd_copy2d_f64(int64_t m, int64_t n, const double* A, int64_t lda, double* B, int64_t ldb, hyd_stream_t s) {
    return hyd::copy2d_local<double>(m, n, A, lda, B, ldb, (cudaStream_t)s);
}
int hyd_copy2d_f32(int64_t m, int64_t n, const float* A, int64_t lda, float* B, int64_t ldb, hyd_stream_t s) {
    return hyd::copy2d_local<float>(m, n, A, lda, B, ldb, (cudaStream_t)s);
}
int hyd_axpy2d_f64(int64_t m, int64_t n, double alpha, const double* A, int64_t lda, double* B, int64_t ldb, hyd_stream_t s) {
    return hyd::axpy2d_local<double>(m, n, alpha, A, lda, B, ldb, (cudaStream_t)s);
}
int hyd_axpy2d_f32(int64_t m, int64_t n, float alpha, const float* A, int64_t lda, float* B, int64_t ldb, hyd_stream_t s) {
    return hyd::axpy2d_local<float>(m, n, alpha, A, lda, B, ldb, (cudaStream_t)s);
}
int hyd_scale2d_f64(int64_t m, int64_t n, double alpha, double* A, int64_t lda, hyd_stream_t s) {
    return hyd::scale_local<double>(m, n, alpha, A, lda, (cudaStream_t)s);
}
int hyd_scale2d_f32(int64_t m, int64_t n, float alpha, float* A, int64_t lda, hyd_stream_t s) {
    return hyd::scale_local<float>(m, n, alpha, A, lda, (cudaStream_t)s);
}

// ---- half-precision storage (fp16 / bf16), fp32 accumulation: SUMMA-C and the 2-byte redistributions ----
int hyd_gemm_f16(hyd_grid* g, int oA, int oB, int alg, float alpha, const hyd_dm* A, const hyd_dm* B, float beta, const hyd_dm* C,
                 int64_t blocksize, hyd_stream_t stream) {
    const hyd::GridShape& sh = g->g->shape;
    if (check_dm(sh, A, "hyd_gemm_f16(A)") || check_dm(sh, B, "hyd_gemm_f16(B)") || check_dm(sh, C, "hyd_gemm_f16(C)")) return 1;
    return hyd::dist_gemm_half(*g->g, 0, oA, oB, alg, alpha, to_dm(A), to_dm(B), beta, to_dm(C), blocksize, (cudaStream_t)stream);
}
int hyd_gemm_bf16(hyd_grid* g, int oA, int oB, int alg, float alpha, const hyd_dm* A, const hyd_dm* B, float beta, const hyd_dm* C,
                  int64_t blocksize, hyd_stream_t stream) {
    const hyd::GridShape& sh = g->g->shape;
    if (check_dm(sh, A, "hyd_gemm_bf16(A)") || check_dm(sh, B, "hyd_gemm_bf16(B)") || check_dm(sh, C, "hyd_gemm_bf16(C)")) return 1;
    return hyd::dist_gemm_half(*g->g, 1, oA, oB, alg, alpha, to_dm(A), to_dm(B), beta, to_dm(C), blocksize, (cudaStream_t)stream);
}
int hyd_redist_b16(hyd_grid* g, const hyd_dm* src, const hyd_dm* dst, int transpose, hyd_stream_t stream) {
    const hyd::GridShape& sh = g->g->shape;
    if (check_dm(sh, src, "hyd_redist_b16(src)") || check_dm(sh, dst, "hyd_redist_b16(dst)")) return 1;
    return hyd::redistribute<hyd::h16>(*g->g, to_dm(src), to_dm(dst), transpose != 0, 3, (cudaStream_t)stream);
}

// ---- triangular rank-k updates on the SUMMA-C engine, and the trapezoid level-1 kernels around them ----
#define HYD_TRRK(SFX, T)                                                                                                        \
    int hyd_trrk_##SFX(hyd_grid* g, int uplo, int oA, int oB, T alpha, const hyd_dm* A, const hyd_dm* B, T beta, const hyd_dm* C, \
                       int64_t blocksize, hyd_stream_t stream) {                                                                \
        const hyd::GridShape& sh = g->g->shape;                                                                                 \
        if (check_dm(sh, A, "hyd_trrk(A)") || check_dm(sh, B, "hyd_trrk(B)") || check_dm(sh, C, "hyd_trrk(C)")) return 1;         \
        return hyd::dist_trrk<T>(*g->g, uplo, oA, oB, alpha, to_dm(A), to_dm(B), beta, to_dm(C), blocksize, (cudaStream_t)stream); \
    }                                                                                                                           \
    int hyd_syrk_##SFX(hyd_grid* g, int uplo, int orient, T alpha, const hyd_dm* A, T beta, const hyd_dm* C, int64_t blocksize,   \
                       hyd_stream_t stream) {                                                                                   \
        const hyd::GridShape& sh = g->g->shape;                                                                                 \
        if (check_dm(sh, A, "hyd_syrk(A)") || check_dm(sh, C, "hyd_syrk(C)")) return 1;                                         \
        /* NORMAL: C := alpha A A^T + beta C; TRANSPOSE / ADJOINT (real): C := alpha A^T A + beta C */                         \
        return hyd::dist_trrk<T>(*g->g, uplo, orient ? 1 : 0, orient ? 0 : 1, alpha, to_dm(A), to_dm(A), beta, to_dm(C),         \
                                 blocksize, (cudaStream_t)stream);                                                              \
    }                                                                                                                           \
    int hyd_scale_trapezoid_##SFX(T alpha, int uplo, int64_t m, int64_t n, T* A, int64_t lda, int64_t offset, hyd_stream_t s) {  \
        if (uplo != 0 && uplo != 1) return set_error("hyd_scale_trapezoid: uplo must be LOWER (0) or UPPER (1)");               \
        if (alpha == T(1)) return 0;                                                                                            \
        return hyd::trapezoid_local<T>(2, uplo, m, n, 0, 1, 0, 1, offset, alpha, nullptr, 0, A, lda, (cudaStream_t)s);            \
    }                                                                                                                           \
    int hyd_axpy_trapezoid_##SFX(int uplo, T alpha, int64_t m, int64_t n, const T* X, int64_t ldx, T* Y, int64_t ldy,            \
                                 int64_t offset, hyd_stream_t s) {                                                              \
        if (uplo != 0 && uplo != 1) return set_error("hyd_axpy_trapezoid: uplo must be LOWER (0) or UPPER (1)");                \
        return hyd::trapezoid_local<T>(1, uplo, m, n, 0, 1, 0, 1, offset, alpha, X, ldx, Y, ldy, (cudaStream_t)s);               \
    }                                                                                                                           \
    int hyd_make_trapezoidal_##SFX(int uplo, int64_t m, int64_t n, T* A, int64_t lda, int64_t offset, hyd_stream_t s) {          \
        if (uplo != 0 && uplo != 1) return set_error("hyd_make_trapezoidal: uplo must be LOWER (0) or UPPER (1)");              \
        return hyd::trapezoid_local<T>(3, uplo, m, n, 0, 1, 0, 1, offset, T(0), nullptr, 0, A, lda, (cudaStream_t)s);            \
    }                                                                                                                           \
    int hyd_scale_trapezoid_dist_##SFX(hyd_grid* g, T alpha, int uplo, const hyd_dm* A, int64_t offset, hyd_stream_t s) {        \
        if (check_dm(g->g->shape, A, "hyd_scale_trapezoid_dist")) return 1;                                                     \
        if (alpha == T(1)) return 0;                                                                                            \
        return hyd::dist_trapezoid<T>(*g->g, 2, uplo, alpha, nullptr, to_dm(A), offset, (cudaStream_t)s);                       \
    }                                                                                                                           \
    int hyd_axpy_trapezoid_dist_##SFX(hyd_grid* g, int uplo, T alpha, const hyd_dm* X, const hyd_dm* Y, int64_t offset,          \
                                      hyd_stream_t s) {                                                                         \
        if (check_dm(g->g->shape, X, "hyd_axpy_trapezoid_dist(X)") || check_dm(g->g->shape, Y, "hyd_axpy_trapezoid_dist(Y)"))    \
            return 1;                                                                                                           \
        const hyd::DM x = to_dm(X);                                                                                             \
        return hyd::dist_trapezoid<T>(*g->g, 1, uplo, alpha, &x, to_dm(Y), offset, (cudaStream_t)s);                            \
    }                                                                                                                           \
    int hyd_symm_##SFX(hyd_grid* g, int side, int uplo, T alpha, const hyd_dm* A, const hyd_dm* B, T beta, const hyd_dm* C,        \
                       int64_t blocksize, hyd_stream_t stream) {                                                                \
        const hyd::GridShape& sh = g->g->shape;                                                                                 \
        if (check_dm(sh, A, "hyd_symm(A)") || check_dm(sh, B, "hyd_symm(B)") || check_dm(sh, C, "hyd_symm(C)")) return 1;         \
        return hyd::dist_symm<T>(*g->g, side, uplo, alpha, to_dm(A), to_dm(B), beta, to_dm(C), blocksize, (cudaStream_t)stream);  \
    }                                                                                                                           \
    int hyd_make_symmetric_dist_##SFX(hyd_grid* g, int uplo, const hyd_dm* A, hyd_stream_t s) {                                  \
        if (check_dm(g->g->shape, A, "hyd_make_symmetric_dist")) return 1;                                                      \
        return hyd::dist_make_symmetric<T>(*g->g, uplo, to_dm(A), (cudaStream_t)s);                                             \
    }                                                                                                                           \
    int hyd_make_trapezoidal_dist_##SFX(hyd_grid* g, int uplo, const hyd_dm* A, int64_t offset, hyd_stream_t s) {                \
        if (check_dm(g->g->shape, A, "hyd_make_trapezoidal_dist")) return 1;                                                    \
        return hyd::dist_trapezoid<T>(*g->g, 3, uplo, T(0), nullptr, to_dm(A), offset, (cudaStream_t)s);                        \
    }
HYD_TRRK(f64, double)
HYD_TRRK(f32, float)
#undef HYD_TRRK

// MakeMultiSync(master, others...) (include/hydrogen/SyncInfo.hpp:136-186): at entry the master stream waits for everything
// enqueued so far on the other streams; at exit the other streams wait for the master.  Events come from a small per-thread
// cache (never destroyed while recorded work may be pending; cudaEventDisableTiming).
static cudaEvent_t multisync_event(int i) {
    static thread_local cudaEvent_t ev[16] = {};
    static thread_local int dev_of[16];
    int dev = 0;
    cudaGetDevice(&dev);
    i &= 15;
    if (ev[i] && dev_of[i] != dev) { cudaEventDestroy(ev[i]); ev[i] = nullptr; }
    if (!ev[i]) { cudaEventCreateWithFlags(&ev[i], cudaEventDisableTiming); dev_of[i] = dev; }
    return ev[i];
}
int hyd_multisync_begin(hyd_stream_t master, const hyd_stream_t* others, int n) {
    if (n < 0 || (n > 0 && !others)) return set_error("hyd_multisync_begin: bad arguments");
    cudaStream_t m = (cudaStream_t)master;
    for (int i = 0; i < n; ++i) {
        cudaStream_t o = (cudaStream_t)others[i];
        if (o == m) continue;
        bool seen = false;
        for (int j = 0; j < i; ++j) seen = seen || others[j] == others[i];
        if (seen) continue;
        cudaEvent_t e = multisync_event(i);
        if (hyd::check_cuda(cudaEventRecord(e, o), "hyd_multisync_begin: record") ||
            hyd::check_cuda(cudaStreamWaitEvent(m, e, 0), "hyd_multisync_begin: wait"))
            return 1;
    }
    return 0;
}
int hyd_multisync_end(hyd_stream_t master, const hyd_stream_t* others, int n) {
    if (n < 0 || (n > 0 && !others)) return set_error("hyd_multisync_end: bad arguments");
    cudaStream_t m = (cudaStream_t)master;
    bool any = false;
    for (int i = 0; i < n; ++i) any = any || (cudaStream_t)others[i] != m;
    if (!any) return 0;
    cudaEvent_t e = multisync_event(15);
    if (hyd::check_cuda(cudaEventRecord(e, m), "hyd_multisync_end: record")) return 1;
    for (int i = 0; i < n; ++i) {
        cudaStream_t o = (cudaStream_t)others[i];
        if (o == m) continue;
        bool seen = false;
        for (int j = 0; j < i; ++j) seen = seen || others[j] == others[i];
        if (seen) continue;
        if (hyd::check_cuda(cudaStreamWaitEvent(o, e, 0), "hyd_multisync_end: wait")) return 1;
    }
    return 0;
}

static int copy_kind(int kind, cudaMemcpyKind* out, const char* who) {
    if (kind == 0) *out = cudaMemcpyHostToDevice;
    else if (kind == 1) *out = cudaMemcpyDeviceToHost;
    else if (kind == 2) *out = cudaMemcpyDeviceToDevice;
    else return set_error("%s: kind must be 0 (H2D), 1 (D2H) or 2 (D2D)", who);
    return 0;
}
int hyd_memcopy2d_async(void* dst, int64_t dst_ldim, const void* src, int64_t src_ldim, int64_t height, int64_t width,
                        int elem_size, int kind, hyd_stream_t stream) {
    const char* who = "hyd_memcopy2d_async";
    cudaMemcpyKind k = cudaMemcpyDefault;
    if (copy_kind(kind, &k, who)) return 1;
    if (elem_size != 4 && elem_size != 8) return set_error("%s: elem_size must be 4 or 8", who);
    if (height < 0 || width < 0) return set_error("%s: negative dimension", who);
    if (height == 0 || width == 0) return 0;
    if (!dst || !src) return set_error("%s: null buffer", who);
    if (dst_ldim < height || src_ldim < height) return set_error("%s: leading dimension too small", who);
    return hyd::check_cuda(cudaMemcpy2DAsync(dst, (size_t)dst_ldim * elem_size, src, (size_t)src_ldim * elem_size,
                                             (size_t)height * elem_size, (size_t)width, k, (cudaStream_t)stream), who);
}
int hyd_memcopy1d_async(void* dst, const void* src, int64_t count, int elem_size, int kind, hyd_stream_t stream) {
    const char* who = "hyd_memcopy1d_async";
    cudaMemcpyKind k = cudaMemcpyDefault;
    if (copy_kind(kind, &k, who)) return 1;
    if (elem_size != 4 && elem_size != 8) return set_error("%s: elem_size must be 4 or 8", who);
    if (count < 0) return set_error("%s: negative count", who);
    if (count == 0) return 0;
    if (!dst || !src) return set_error("%s: null buffer", who);
    return hyd::check_cuda(cudaMemcpyAsync(dst, src, (size_t)count * elem_size, k, (cudaStream_t)stream), who);
}

int hyd_gemm_host_f64(hyd_grid* g, int oA, int oB, int alg, double alpha, const hyd_dm* A, const hyd_dm* B, double beta,
                      const hyd_dm* C, int64_t blocksize, hyd_stream_t stream) {
    return gemm_host<double>(g, oA, oB, alg, alpha, A, B, beta, C, blocksize, (cudaStream_t)stream);
}
int hyd_gemm_host_f32(hyd_grid* g, int oA, int oB, int alg, float alpha, const hyd_dm* A, const hyd_dm* B, float beta,
                      const hyd_dm* C, int64_t blocksize, hyd_stream_t stream) {
    return gemm_host<float>(g, oA, oB, alg, alpha, A, B, beta, C, blocksize, (cudaStream_t)stream);
}

}  // extern "C"
