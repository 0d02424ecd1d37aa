// Multi-GPU layer of the B200 sunsky path (include/sunsky_b200.h, "multi-GPU" section). The reference has no
// multi-device code at all (no NCCL / MPI anywhere under src/ or include/); BASELINE.json's north
// star asks for ray / pixel batches sharded over the GPUs of one 8 x B200 box with no inter-GPU traffic except an
// optional gather of the output images. Everything here sits on top of the single-GPU C ABI (ssk_api.cu):
//
//   * sharding arithmetic (contiguous ray ranges, blocked / round-robin frames)
//   * ssk_sweep_render: BASELINE configs[3] / [4], many emitters -> one env map each, re-parametrised in batches
//   * frames gathered on one GPU in three ways:
//       - peer stores (the product path): the destination buffer lives on the root GPU and is mapped into the
//         other processes (CUDA IPC) or is simply addressable (one process, peer access); envmap_kernel writes
//         its pixels straight into it over NVLink while it computes -- compute and "collective" are one kernel,
//         nothing is staged, nothing is copied afterwards
//       - ncclSend / ncclRecv groups on a second stream, double-buffered against the next chunk's kernels
//         (libnccl is dlopen'ed: the library has no link-time dependency on it)
//   * ssk_group_*: one process driving N GPUs with one host thread per GPU (the "C++ host code, no PyTorch"
//     driver of the north star); bench.py's torchrun ranks use the same entry points one GPU each.
#include "../../include/sunsky_b200.h"
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <algorithm>
#include <cstring>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

namespace {

thread_local std::string g_merr;
int mfail(int code, const std::string &m) { g_merr = m; return code; }
#define MCUDA(call)                                                                                \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess)                                                                     \
            return mfail(SSK_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_));        \
    } while (0)

struct DevScope {
    int prev = -1; bool restore = false;
    int enter(int device) {
        if (cudaGetDevice(&prev) != cudaSuccess) { cudaGetLastError(); prev = -1; }
        if (prev != device) { MCUDA(cudaSetDevice(device)); restore = prev >= 0; }
        return SSK_OK;
    }
    ~DevScope() { if (restore) cudaSetDevice(prev); }
};

// ---- NCCL through dlopen --------------------------------------------------------------------------------------
typedef struct ncclComm *ncclComm_t;
struct ncclUniqueId_ { char internal[128]; };
enum { ncclSuccess_ = 0, ncclFloat_ = 7 };
struct NcclApi {
    void *h = nullptr;
    int (*GetUniqueId)(ncclUniqueId_ *) = nullptr;
    int (*CommInitRank)(ncclComm_t *, int, ncclUniqueId_, int) = nullptr;
    int (*CommDestroy)(ncclComm_t) = nullptr;
    int (*Send)(const void *, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*Recv)(void *, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
    int (*GetVersion)(int *) = nullptr;
};
NcclApi *nccl() {
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        // a copy already in the process (torch's bundled libnccl.so.2) is reused: same soname
        for (const char *name : { "libnccl.so.2", "libnccl.so" }) {
            api.h = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
            if (api.h) break;
        }
        if (!api.h) return;
        auto sym = [&](const char *n) { return dlsym(api.h, n); };
        api.GetUniqueId = (decltype(api.GetUniqueId)) sym("ncclGetUniqueId");
        api.CommInitRank = (decltype(api.CommInitRank)) sym("ncclCommInitRank");
        api.CommDestroy = (decltype(api.CommDestroy)) sym("ncclCommDestroy");
        api.Send = (decltype(api.Send)) sym("ncclSend");
        api.Recv = (decltype(api.Recv)) sym("ncclRecv");
        api.GroupStart = (decltype(api.GroupStart)) sym("ncclGroupStart");
        api.GroupEnd = (decltype(api.GroupEnd)) sym("ncclGroupEnd");
        api.GetErrorString = (decltype(api.GetErrorString)) sym("ncclGetErrorString");
        api.GetVersion = (decltype(api.GetVersion)) sym("ncclGetVersion");
        if (!api.GetUniqueId || !api.CommInitRank || !api.Send || !api.Recv || !api.GroupStart || !api.GroupEnd) api.h = nullptr;
    });
    return api.h ? &api : nullptr;
}
#define MNCCL(call)                                                                                \
    do {                                                                                           \
        int r_ = (call);                                                                           \
        if (r_ != ncclSuccess_)                                                                    \
            return mfail(SSK_ERR_CUDA, std::string(#call) + ": " +                                 \
                         (nccl()->GetErrorString ? nccl()->GetErrorString(r_) : "NCCL error"));    \
    } while (0)

} // namespace

struct ssk_comm {
    ncclComm_t comm = nullptr;
    int rank = 0, world = 1, device = 0;
    cudaStream_t stream = nullptr;          // the gather stream (second stream of the double-buffered sweep)
};

extern "C" {

const char *ssk_multi_last_error(void) { return g_merr.c_str(); }

// ---- sharding ----------------------------------------------------------------------------------
int ssk_shard_ray_range(size_t n, int rank, int world, size_t align, size_t *begin, size_t *end) {
    if (world < 1 || rank < 0 || rank >= world || align < 1 || !begin || !end)
        return mfail(SSK_ERR_INVALID, "bad rank / world size / alignment");
    const size_t units = n / align, base = units / (size_t) world, extra = units % (size_t) world;
    const size_t r = (size_t) rank;
    const size_t b = r * base + std::min(r, extra), e = b + base + (r < extra ? 1 : 0);
    *begin = b * align;
    *end = rank == world - 1 ? n : e * align;
    return SSK_OK;
}

int ssk_shard_frames(int n_frames, int rank, int world, int mode, int *frames, int *count) {
    if (world < 1 || rank < 0 || rank >= world || n_frames < 0 || !count) return mfail(SSK_ERR_INVALID, "bad rank / world size");
    int k = 0;
    if (mode == SSK_SHARD_ROUND_ROBIN) {
        for (int f = rank; f < n_frames; f += world) { if (frames) frames[k] = f; ++k; }
    } else if (mode == SSK_SHARD_BLOCKED) {
        size_t b, e;
        ssk_shard_ray_range((size_t) n_frames, rank, world, 1, &b, &e);
        for (size_t f = b; f < e; ++f) { if (frames) frames[k] = (int) f; ++k; }
    } else return mfail(SSK_ERR_INVALID, "mode must be SSK_SHARD_BLOCKED or SSK_SHARD_ROUND_ROBIN");
    *count = k;
    return SSK_OK;
}

// ---- peer-mapped frame buffers --------------------------------------------------------------------------------
int ssk_peer_alloc(int device, size_t bytes, void **ptr, unsigned char handle[SSK_PEER_HANDLE_BYTES]) {
    if (!ptr || !bytes) return mfail(SSK_ERR_INVALID, "null argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == SSK_PEER_HANDLE_BYTES, "handle size");
    DevScope ds; if (int rc = ds.enter(device)) return rc;
    MCUDA(cudaMalloc(ptr, bytes));
    if (handle) {
        cudaIpcMemHandle_t h;
        cudaError_t e = cudaIpcGetMemHandle(&h, *ptr);
        if (e != cudaSuccess) { cudaFree(*ptr); *ptr = nullptr; return mfail(SSK_ERR_CUDA, std::string("cudaIpcGetMemHandle: ") + cudaGetErrorString(e)); }
        memcpy(handle, &h, sizeof(h));
    }
    return SSK_OK;
}
int ssk_peer_open(int device, const unsigned char handle[SSK_PEER_HANDLE_BYTES], void **ptr) {
    if (!ptr || !handle) return mfail(SSK_ERR_INVALID, "null argument");
    DevScope ds; if (int rc = ds.enter(device)) return rc;
    cudaIpcMemHandle_t h; memcpy(&h, handle, sizeof(h));
    MCUDA(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return SSK_OK;
}
int ssk_peer_close(int device, void *ptr) {
    DevScope ds; if (int rc = ds.enter(device)) return rc;
    MCUDA(cudaIpcCloseMemHandle(ptr));
    return SSK_OK;
}
int ssk_peer_free(int device, void *ptr) {
    DevScope ds; if (int rc = ds.enter(device)) return rc;
    MCUDA(cudaFree(ptr));
    return SSK_OK;
}

// ---- sweeps: many emitters, one env map each -----------------------------------------------------------------
namespace {
struct SweepPool {                      // contexts reused across calls: one pool per (device, mode, arith)
    int device, mode, arith;
    std::string dir;
    std::vector<ssk_ctx *> ctx;
    std::mutex mu;                      // one sweep at a time per pool; different devices run concurrently
};
std::mutex g_pool_mu;                   // guards the list only
std::vector<std::unique_ptr<SweepPool>> g_pools;
}

int ssk_sweep_render(const ssk_params *frames, int n_frames, const char *dataset_dir, int device, int mode,
                     int width, int height, const float *wavelengths4, float *const *out, int chunk, void *stream) {
    if (n_frames < 0 || (n_frames && (!frames || !out)) || !dataset_dir || width <= 0 || height <= 0)
        return mfail(SSK_ERR_INVALID, "bad sweep arguments");
    if (n_frames == 0) return SSK_OK;
    if (mode == SSK_MODE_SPECTRAL && !wavelengths4) return mfail(SSK_ERR_INVALID, "spectral sweeps need 4 wavelengths per frame");
    if (chunk < 1) chunk = 8;
    chunk = std::min(chunk, n_frames);
    SweepPool *pool = nullptr;
    {
        std::lock_guard<std::mutex> lock(g_pool_mu);
        for (auto &p : g_pools)
            if (p->device == device && p->mode == mode && p->arith == frames[0].arith && p->dir == dataset_dir) pool = p.get();
        if (!pool) {
            g_pools.emplace_back(new SweepPool());
            pool = g_pools.back().get();
            pool->device = device; pool->mode = mode; pool->arith = frames[0].arith; pool->dir = dataset_dir;
        }
    }
    std::lock_guard<std::mutex> lock(pool->mu);
    for (int s = 0; s < n_frames; s += chunk) {
        const int m = std::min(chunk, n_frames - s);
        // parameters_changed for the whole chunk: the contexts' previous launches may still be queued -- updates
        // write the next block of each context's ring (ssk_api.cu), so nothing here waits for the device except
        // the chunk's own setup quadrature (its stream has the highest priority and overlaps the queued maps)
        const int have = (int) pool->ctx.size();
        const int upd = std::min(have, m);
        if (upd > 0) {
            // a frame is a full property set: constructor semantics, i.e. the sun direction is normalised
            // (parameters_changed() alone would take it as stored)
            std::vector<ssk_params> fp(frames + s, frames + s + upd);
            for (auto &q : fp)
                if (!q.use_record && !(q.flags & SSK_FLAG_SUN_DIR_NORMALIZED)) ssk_normalize_direction(q.sun_direction, q.sun_direction);
            if (int rc = ssk_update_batch(pool->ctx.data(), fp.data(), upd)) return mfail(rc, ssk_last_error());
        }
        for (int i = upd; i < m; ++i) {
            ssk_ctx *c = nullptr;
            if (int rc = ssk_create(frames + s + i, dataset_dir, device, mode, &c)) return mfail(rc, ssk_last_error());
            pool->ctx.push_back(c);
        }
        for (int i = 0; i < m; ++i) {
            ssk_ctx *c = pool->ctx[i];
            if (int rc = ssk_envmap_device(&c, 1, width, height, wavelengths4 ? wavelengths4 + 4 * (size_t) (s + i) : nullptr,
                                           out[s + i], stream))
                return mfail(rc, ssk_last_error());
        }
    }
    return SSK_OK;
}

void ssk_sweep_release(void) {
    std::lock_guard<std::mutex> lock(g_pool_mu);
    for (auto &p : g_pools) for (ssk_ctx *c : p->ctx) ssk_destroy(c);
    g_pools.clear();
}

// ---- NCCL: frames to one rank --------------------------------------------------------------------------------
int ssk_comm_unique_id(unsigned char id[SSK_COMM_ID_BYTES]) {
    if (!id) return mfail(SSK_ERR_INVALID, "null argument");
    if (!nccl()) return mfail(SSK_ERR_UNSUPPORTED, "libnccl.so.2 could not be loaded");
    ncclUniqueId_ u;
    MNCCL(nccl()->GetUniqueId(&u));
    memcpy(id, &u, sizeof(u));
    return SSK_OK;
}
int ssk_comm_create(const unsigned char id[SSK_COMM_ID_BYTES], int rank, int world, int device, ssk_comm **out) {
    if (!id || !out || world < 1 || rank < 0 || rank >= world) return mfail(SSK_ERR_INVALID, "bad communicator arguments");
    if (!nccl()) return mfail(SSK_ERR_UNSUPPORTED, "libnccl.so.2 could not be loaded");
    DevScope ds; if (int rc = ds.enter(device)) return rc;
    ssk_comm *c = new ssk_comm();
    c->rank = rank; c->world = world; c->device = device;
    ncclUniqueId_ u; memcpy(&u, id, sizeof(u));
    int r = nccl()->CommInitRank(&c->comm, world, u, rank);
    if (r != ncclSuccess_) { delete c; return mfail(SSK_ERR_CUDA, std::string("ncclCommInitRank: ") + (nccl()->GetErrorString ? nccl()->GetErrorString(r) : "?")); }
    cudaError_t e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
    if (e != cudaSuccess) { nccl()->CommDestroy(c->comm); delete c; return mfail(SSK_ERR_CUDA, cudaGetErrorString(e)); }
    *out = c;
    return SSK_OK;
}
void ssk_comm_destroy(ssk_comm *c) {
    if (!c) return;
    DevScope ds; ds.enter(c->device);
    if (c->stream) { cudaStreamSynchronize(c->stream); cudaStreamDestroy(c->stream); }
    if (c->comm && nccl() && nccl()->CommDestroy) nccl()->CommDestroy(c->comm);
    delete c;
}
int ssk_comm_version(void) { int v = 0; if (nccl() && nccl()->GetVersion) nccl()->GetVersion(&v); return v; }

/// Frames `first .. first + count - 1` of every rank's local list (ssk_shard_frames order) to `root`, straight into
/// their final positions of `full` ([n_frames][frame_floats] on the root). One ncclSend per local frame, one ncclRecv
/// per remote frame, all in one group (NCCL matches the operations of a pair of ranks in order); the root's own
/// frames are device-to-device copies. Ranks own different frame counts when n_frames % world != 0: a rank whose
/// list is exhausted simply posts nothing. Enqueued on `stream`; returns without synchronising.
int ssk_comm_gather_frames(ssk_comm *c, const float *local, size_t frame_floats, int n_frames, int mode, int first, int count,
                           float *full, int root, void *stream) {
    if (!c || frame_floats == 0 || n_frames < 0 || first < 0 || count < 0 || root < 0 || root >= c->world)
        return mfail(SSK_ERR_INVALID, "bad gather arguments");
    if (c->rank == root && !full) return mfail(SSK_ERR_INVALID, "the root needs the destination buffer");
    DevScope ds; if (int rc = ds.enter(c->device)) return rc;
    cudaStream_t s = stream ? (cudaStream_t) stream : c->stream;
    std::vector<int> idx((size_t) std::max(n_frames, 1));
    int mine = 0;
    if (int rc = ssk_shard_frames(n_frames, c->rank, c->world, mode, idx.data(), &mine)) return rc;
    const int lo = std::min(first, mine), hi = std::min(first + count, mine);
    if (hi > lo && !local) return mfail(SSK_ERR_INVALID, "null local frames");
    if (c->rank != root) {
        if (hi > lo) {
            MNCCL(nccl()->GroupStart());
            for (int k = lo; k < hi; ++k)
                MNCCL(nccl()->Send(local + (size_t) (k - first) * frame_floats, frame_floats, ncclFloat_, root, c->comm, s));
            MNCCL(nccl()->GroupEnd());
        }
        return SSK_OK;
    }
    for (int k = lo; k < hi; ++k)
        MCUDA(cudaMemcpyAsync(full + (size_t) idx[k] * frame_floats, local + (size_t) (k - first) * frame_floats,
                              frame_floats * sizeof(float), cudaMemcpyDeviceToDevice, s));
    MNCCL(nccl()->GroupStart());
    std::vector<int> theirs((size_t) std::max(n_frames, 1));
    for (int r = 0; r < c->world; ++r) {
        if (r == root) continue;
        int cnt = 0;
        ssk_shard_frames(n_frames, r, c->world, mode, theirs.data(), &cnt);
        for (int k = std::min(first, cnt); k < std::min(first + count, cnt); ++k)
            MNCCL(nccl()->Recv(full + (size_t) theirs[k] * frame_floats, frame_floats, ncclFloat_, r, c->comm, s));
    }
    MNCCL(nccl()->GroupEnd());
    return SSK_OK;
}

/// configs[3] / [4] with the frames collected on `root` through NCCL: this rank renders its frames chunk by chunk
/// into two local staging buffers on `stream`; as soon as a chunk is rendered its frames travel on the
/// communicator's own stream while the next chunk renders into the other buffer.
int ssk_sweep_render_gather_nccl(ssk_comm *c, const ssk_params *frames /* this rank's, shard order */, int n_frames_total,
                                 int shard_mode, const char *dataset_dir, int mode, int width, int height,
                                 const float *wavelengths4, float *staging /* 2 * chunk frames */, float *full, int root,
                                 int chunk, void *stream) {
    if (!c || !staging || chunk < 1) return mfail(SSK_ERR_INVALID, "bad arguments");
    DevScope ds; if (int rc = ds.enter(c->device)) return rc;
    const size_t frame_floats = (size_t) (mode == SSK_MODE_SPECTRAL ? 4 : 3) * width * height;
    int mine = 0, most = 0;
    if (int rc = ssk_shard_frames(n_frames_total, c->rank, c->world, shard_mode, nullptr, &mine)) return rc;
    for (int r = 0; r < c->world; ++r) { int cnt = 0; ssk_shard_frames(n_frames_total, r, c->world, shard_mode, nullptr, &cnt); most = std::max(most, cnt); }
    cudaEvent_t rendered[2], sent[2];
    for (int i = 0; i < 2; ++i) { MCUDA(cudaEventCreateWithFlags(&rendered[i], cudaEventDisableTiming)); MCUDA(cudaEventCreateWithFlags(&sent[i], cudaEventDisableTiming)); }
    int rc = SSK_OK;
    cudaStream_t s = (cudaStream_t) stream;
    for (int first = 0, k = 0; first < most && rc == SSK_OK; first += chunk, ++k) {
        const int slot = k & 1;
        float *buf = staging + (size_t) slot * chunk * frame_floats;
        const int m = std::max(0, std::min(chunk, mine - first));
        if (k >= 2) cudaStreamWaitEvent(s, sent[slot], 0);               // the buffer's previous frames have left
        if (m > 0) {
            std::vector<float *> outs((size_t) m);
            for (int i = 0; i < m; ++i) outs[i] = buf + (size_t) i * frame_floats;
            rc = ssk_sweep_render(frames + first, m, dataset_dir, c->device, mode, width, height,
                                  wavelengths4 ? wavelengths4 + 4 * (size_t) first : nullptr, outs.data(), chunk, s);
            if (rc) break;
        }
        cudaEventRecord(rendered[slot], s);
        cudaStreamWaitEvent(c->stream, rendered[slot], 0);
        rc = ssk_comm_gather_frames(c, buf, frame_floats, n_frames_total, shard_mode, first, chunk, full, root, c->stream);
        cudaEventRecord(sent[slot], c->stream);
    }
    if (rc == SSK_OK) { cudaStreamWaitEvent(s, sent[0], 0); cudaStreamWaitEvent(s, sent[1], 0); }   // `stream` sees the gather done
    for (int i = 0; i < 2; ++i) { cudaEventDestroy(rendered[i]); cudaEventDestroy(sent[i]); }
    return rc;
}

} // extern "C"

// ---- one process, N GPUs ---------------------------------------------------------------------------------------
struct ssk_group {
    std::vector<int> devices;
    std::vector<ssk_ctx *> ctx;             // one emitter context per device, same properties
    int mode = SSK_MODE_RGB;
    std::string dir;
};

extern "C" {

int ssk_group_create(const int *devices, int n_devices, const ssk_params *p, const char *dataset_dir, int mode, ssk_group **out) {
    if (!devices || n_devices < 1 || !p || !dataset_dir || !out) return mfail(SSK_ERR_INVALID, "null argument");
    ssk_group *g = new ssk_group();
    g->mode = mode; g->dir = dataset_dir;
    for (int i = 0; i < n_devices; ++i) {
        ssk_ctx *c = nullptr;
        if (int rc = ssk_create(p, dataset_dir, devices[i], mode, &c)) {
            std::string msg = ssk_last_error();
            for (ssk_ctx *q : g->ctx) ssk_destroy(q);
            delete g;
            return mfail(rc, msg);
        }
        g->devices.push_back(devices[i]); g->ctx.push_back(c);
    }
    // every GPU may store into every other GPU's memory (NVLink / NVSwitch): the gathered sweeps rely on it
    for (int a : g->devices)
        for (int b : g->devices) {
            if (a == b) continue;
            int can = 0;
            cudaDeviceCanAccessPeer(&can, a, b);
            if (!can) continue;
            DevScope ds; ds.enter(a);
            cudaError_t e = cudaDeviceEnablePeerAccess(b, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { /* reported when a gather needs it */ }
            cudaGetLastError();
        }
    *out = g;
    return SSK_OK;
}
void ssk_group_destroy(ssk_group *g) {
    if (!g) return;
    for (ssk_ctx *c : g->ctx) ssk_destroy(c);
    delete g;
}
int ssk_group_size(const ssk_group *g) { return g ? (int) g->devices.size() : 0; }
ssk_ctx *ssk_group_context(const ssk_group *g, int i) { return (g && i >= 0 && i < (int) g->ctx.size()) ? g->ctx[i] : nullptr; }

int ssk_group_update(ssk_group *g, const ssk_params *p) {
    if (!g || !p) return mfail(SSK_ERR_INVALID, "null argument");
    for (ssk_ctx *c : g->ctx) if (int rc = ssk_update(c, p)) return mfail(rc, ssk_last_error());
    return SSK_OK;
}

} // extern "C"

namespace {
/// Runs f(i) for every device of the group on its own host thread; first failure wins
template <typename F> int per_device(const ssk_group *g, F &&f) {
    const int n = (int) g->devices.size();
    std::vector<int> rc((size_t) n, SSK_OK);
    std::vector<std::string> msg((size_t) n);
    std::vector<std::thread> th;
    for (int i = 0; i < n; ++i)
        th.emplace_back([&, i] { rc[i] = f(i); if (rc[i]) msg[i] = ssk_last_error()[0] ? ssk_last_error() : ssk_multi_last_error(); });
    for (auto &t : th) t.join();
    for (int i = 0; i < n; ++i) if (rc[i]) return mfail(rc[i], "device " + std::to_string(g->devices[i]) + ": " + msg[i]);
    return SSK_OK;
}
}

extern "C" {

/// eval over host planes, the batch split into contiguous 4-ray-aligned ranges, one range per GPU, each through
/// ssk_eval_host (H2D + kernel + D2H) on its own host thread
int ssk_group_eval_host(const ssk_group *g, const float *const wi[3], const float *const wav[4], size_t n, float *const out[4]) {
    if (!g || !wi || !out) return mfail(SSK_ERR_INVALID, "null argument");
    const int C = g->mode == SSK_MODE_SPECTRAL ? 4 : 3, W = (int) g->devices.size();
    return per_device(g, [&](int i) -> int {
        size_t b, e;
        ssk_shard_ray_range(n, i, W, 4, &b, &e);
        if (e <= b) return SSK_OK;
        const float *w3[3] = { wi[0] + b, wi[1] + b, wi[2] + b };
        const float *w4[4] = { nullptr, nullptr, nullptr, nullptr };
        if (wav) for (int c = 0; c < 4; ++c) w4[c] = wav[c] ? wav[c] + b : nullptr;
        float *o[4] = { nullptr, nullptr, nullptr, nullptr };
        for (int c = 0; c < C; ++c) o[c] = out[c] + b;
        return ssk_eval_host(g->ctx[i], w3, wav ? w4 : nullptr, e - b, o);
    });
}

int ssk_group_sample_direction_host(const ssk_group *g, const float *const u[2], const float *const wav[4], size_t n,
                                    float *const d[3], float *pdf, float *const weight[4]) {
    if (!g || !u || !d || !pdf || !weight) return mfail(SSK_ERR_INVALID, "null argument");
    const int C = g->mode == SSK_MODE_SPECTRAL ? 4 : 3, W = (int) g->devices.size();
    return per_device(g, [&](int i) -> int {
        size_t b, e;
        ssk_shard_ray_range(n, i, W, 4, &b, &e);
        if (e <= b) return SSK_OK;
        const float *uu[2] = { u[0] + b, u[1] + b };
        const float *w4[4] = { nullptr, nullptr, nullptr, nullptr };
        if (wav) for (int c = 0; c < 4; ++c) w4[c] = wav[c] ? wav[c] + b : nullptr;
        float *dd[3] = { d[0] + b, d[1] + b, d[2] + b };
        float *q[4] = { nullptr, nullptr, nullptr, nullptr };
        for (int c = 0; c < C; ++c) q[c] = weight[c] + b;
        return ssk_sample_direction_host(g->ctx[i], uu, nullptr, wav ? w4 : nullptr, e - b, dd, pdf + b, q, nullptr, nullptr);
    });
}

/// configs[3] / [4] on every GPU of the group. `out` = n_frames destination pointers:
///   gather_device < 0: out[f] must live on the GPU that owns frame f (ssk_shard_frames over the group);
///   gather_device >= 0: all of them live on that one GPU and the other GPUs' kernels store into it over NVLink.
/// Returns after every GPU has finished.
int ssk_group_sweep(const ssk_group *g, const ssk_params *frames, int n_frames, int shard_mode, int width, int height,
                    const float *wavelengths4, float *const *out, int chunk) {
    if (!g || !frames || !out || n_frames < 0) return mfail(SSK_ERR_INVALID, "null argument");
    const int W = (int) g->devices.size();
    return per_device(g, [&](int i) -> int {
        std::vector<int> idx((size_t) std::max(n_frames, 1));
        int mine = 0;
        if (int rc = ssk_shard_frames(n_frames, i, W, shard_mode, idx.data(), &mine)) return rc;
        if (mine == 0) return SSK_OK;
        std::vector<ssk_params> fp((size_t) mine);
        std::vector<float *> o((size_t) mine);
        std::vector<float> wl;
        for (int k = 0; k < mine; ++k) {
            fp[k] = frames[idx[k]]; o[k] = out[idx[k]];
            if (wavelengths4) wl.insert(wl.end(), wavelengths4 + 4 * (size_t) idx[k], wavelengths4 + 4 * (size_t) idx[k] + 4);
        }
        DevScope ds; if (int rc = ds.enter(g->devices[i])) return rc;
        cudaStream_t s;
        MCUDA(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
        int rc = ssk_sweep_render(fp.data(), mine, g->dir.c_str(), g->devices[i], g->mode, width, height,
                                  wavelengths4 ? wl.data() : nullptr, o.data(), chunk, s);
        cudaError_t e = cudaStreamSynchronize(s);
        cudaStreamDestroy(s);
        if (rc == SSK_OK && e != cudaSuccess) return mfail(SSK_ERR_CUDA, cudaGetErrorString(e));
        return rc;
    });
}

} // extern "C"
