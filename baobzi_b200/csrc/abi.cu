// abi.cu -- the C ABI of libbaobzi.so (declared in include/baobzi.h).
//
// Mirrors the reference's src/baobzi.cpp: same exported names, same handle struct, same error
// conventions (NULL + stderr).  Behind the handle sits a bz::Function: the host image of the tree
// (for save / stats / new replicas) plus one device replica per GPU that has been used, each with a
// small pool of execution contexts (stream + chunk buffers) so that concurrent calls on one handle
// from several host threads stay safe, as they are in the reference (eval is const there).
#include "../../include/baobzi.h"
#include "device_tree.hpp"
#include "fit.hpp"
#include "host_tree.hpp"

#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <exception>
#include <memory>
#include <mutex>
#include <algorithm>
#include <string>
#include <thread>
#include <utility>
#include <vector>

namespace bz {

// One in-flight evaluation on one device: a stream, device staging for host-pointer calls, the binned
// workspace, and pinned scratch for the scalar entry point.
struct ExecContext {
    cudaStream_t stream = nullptr;
    // host-pointer path: one chunk in flight per context (several contexts work on one call, see fn_eval_multi)
    double *d_x = nullptr, *d_res = nullptr; // device chunk buffers
    int64_t slot_points = 0;
    double *h_x = nullptr, *h_res = nullptr; // pinned staging for pageable caller memory
    int64_t staged_points = 0;
    BinWorkspace ws;
    double *pinned = nullptr; // scalar eval scratch: x[3], completion word, y[out_dim]
    size_t pinned_doubles = 0;
    unsigned long long scalar_seq = 0;
    bool busy = false;
};

struct Replica {
    DeviceTree dt;
    std::vector<std::unique_ptr<ExecContext>> contexts;
    cudaStream_t shard_stream = nullptr; // baobzi_eval_multi_sharded
    BinWorkspace shard_ws;
    // baobzi_eval_multi_device_dot: per-CTA partial sums, and the chunk of values of the binned / bucketed paths
    double *dot_partial = nullptr, *dot_tmp = nullptr;
    int64_t dot_partial_cap = 0, dot_tmp_cap = 0;
    // shard_ws and the dot buffers are ONE set per device: calls that come in on different streams are chained through
    // this event (wait before the first launch, record after the last), so they never use the workspace concurrently
    cudaEvent_t ws_done = nullptr;
    cudaStream_t ws_last_stream = nullptr;
    bool ws_used = false;
};

enum Mode { kAuto = 0, kDirect = 1, kBinned = 2, kBucketed = 3 };

struct Function {
    HostTree host;
    int host_devices = 1; // GPUs a host-pointer call is spread over (baobzi_set_host_devices / BAOBZI_B200_DEVICES)
    std::mutex mu;
    std::vector<std::unique_ptr<Replica>> replicas;
    std::atomic<int> mode{kAuto};       // baobzi_set_mode may race with evaluations on other threads
    std::atomic<int64_t> chunk{0};
    std::atomic<int64_t> launches{0};
    std::mutex prof_mu;
    std::shared_ptr<StageProfile> profile; // non-null while baobzi_profile_begin .. _end (one profiled caller at a time)
};

namespace {

class DeviceGuard {
  public:
    explicit DeviceGuard(int dev) {
        cudaGetDevice(&prev_);
        if (dev != prev_)
            cudaSetDevice(dev);
        cur_ = dev;
    }
    ~DeviceGuard() {
        if (cur_ != prev_)
            cudaSetDevice(prev_);
    }

  private:
    int prev_ = 0, cur_ = 0;
};

bool report(cudaError_t e, const char *what) {
    if (e == cudaSuccess)
        return false;
    std::fprintf(stderr, "Baobzi (b200) CUDA error in %s: %s\n", what, cudaGetErrorString(e));
    return true;
}

// Points per binned chunk.  Large enough that every leaf fills many warp tiles (padding waste ~ tile / points
// per leaf), small enough that the chunk's keys / slots mostly stay in L2: 2048 points per leaf, clamped.
int64_t default_chunk(const DeviceTree &dt) {
    if (const char *s = std::getenv("BAOBZI_B200_CHUNK")) {
        long long v = std::atoll(s);
        if (v > 0)
            return v;
    }
    int64_t c = (int64_t)dt.view.n_leaves * 2048;
    // 2D: the evaluation is bound by the 16-byte point gather, so the chunk's points must stay in L2 (4 Mi points =
    // 64 MB); 3D is FP64 bound and gathers out of DRAM without loss, so padding waste decides
    // (3D trees with few leaves: measured on the order-12 benchmark tree, 2 Mi -> 8 Mi points per chunk takes the sort
    // from 22.6 to 11.8 ms per 1e9 points and the evaluation from 0.69 to 0.73 of the FP64 peak)
    const int64_t lo = dt.view.dim == 2 ? int64_t(1) << 21 : int64_t(1) << 23;
    const int64_t hi = dt.view.dim == 2 ? int64_t(1) << 22 : int64_t(1) << 24;
    return c < lo ? lo : (c > hi ? hi : c);
}

int env_mode() {
    if (const char *s = std::getenv("BAOBZI_B200_MODE")) {
        if (!std::strcmp(s, "direct"))
            return kDirect;
        if (!std::strcmp(s, "binned"))
            return kBinned;
        if (!std::strcmp(s, "bucketed"))
            return kBucketed;
    }
    return kAuto;
}

// replica on `dev`, created on first use
Replica *get_replica(Function *f, int dev) {
    std::lock_guard<std::mutex> lock(f->mu);
    for (auto &r : f->replicas)
        if (r->dt.device == dev)
            return r.get();
    DeviceGuard g(dev);
    auto r = std::make_unique<Replica>();
    int rc = upload_tree(f->host, r->dt);
    if (rc != 0) {
        report((cudaError_t)rc, "tree upload");
        return nullptr;
    }
    f->replicas.push_back(std::move(r));
    return f->replicas.back().get();
}

// caller holds f->mu
void workspace_begin(Replica *r, cudaStream_t stream) {
    if (r->ws_used && r->ws_last_stream != stream && r->ws_done)
        cudaStreamWaitEvent(stream, r->ws_done, 0);
}

void workspace_end(Replica *r, cudaStream_t stream) {
    if (!r->ws_done)
        cudaEventCreateWithFlags(&r->ws_done, cudaEventDisableTiming);
    if (r->ws_done)
        cudaEventRecord(r->ws_done, stream);
    r->ws_last_stream = stream;
    r->ws_used = true;
}

ExecContext *acquire(Function *f, Replica *r) {
    std::lock_guard<std::mutex> lock(f->mu);
    for (auto &c : r->contexts)
        if (!c->busy) {
            c->busy = true;
            return c.get();
        }
    auto c = std::make_unique<ExecContext>();
    if (report(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking), "stream create"))
        return nullptr;
    c->busy = true;
    r->contexts.push_back(std::move(c));
    return r->contexts.back().get();
}

void release(Function *f, ExecContext *c) {
    std::lock_guard<std::mutex> lock(f->mu);
    c->busy = false;
}

void destroy_context(ExecContext &c) {
    bz_free(c.d_x);
    bz_free(c.d_res);
    if (c.h_x)
        cudaFreeHost(c.h_x);
    if (c.h_res)
        cudaFreeHost(c.h_res);
    free_workspace(c.ws);
    if (c.pinned)
        cudaFreeHost(c.pinned);
    if (c.stream)
        cudaStreamDestroy(c.stream);
}

// The reference API asks nothing of `x` beyond being a double array, so a 2D point array may start on an odd
// double (a view one element into a buffer).  The grouped 2D kernels gather a point as ONE 16-byte item
// (cp.async 16 in bin_eval2d_kernel); such input takes the direct kernel, whose loads fall back to two 8-byte ones
// (lookup.cuh load_point<2>).  1D / 3D kernels only ever issue 8-byte accesses to x.
bool misaligned_2d(const DeviceTree &dt, const double *d_x) {
    return dt.view.dim == 2 && (reinterpret_cast<uintptr_t>(d_x) & 15u) != 0;
}

bool use_binned(const Function *f, const DeviceTree &dt, int64_t n) {
    if (!binned_supported(dt.view.dim, dt.view.order) || !binned_fits(dt))
        return false;
    const int mode = f->mode.load();
    if (mode == kDirect || mode == kBucketed)
        return false;
    if (mode == kBinned)
        return true;
    // auto: worth it when the per-point coefficient gather (order^dim doubles) dwarfs the point itself and
    // there are enough points per leaf to fill warp tiles
    const int64_t block_bytes = 8 * (int64_t)std::pow((double)dt.view.order, dt.view.dim);
    if (dt.view.dim == 2) // every lane gathering its own block is bound by the L1 tag rate at any block size
        return n >= 64 * (int64_t)dt.view.n_leaves && n >= (1 << 19); // measured crossover (C1 tree): 0.25-1 M points
    // 3D trees far larger than L2 with so many leaves that a chunk holds only a handful of points per leaf: the evaluation
    // tiles are mostly padding and the four-kernel sort scans a huge histogram; the bucketed path is ahead there (order 6,
    // 997 K leaves, 1.8 GB: 5.0e9 against 3.7e9 evals/s; at 249 K leaves / 449 MB the binned path still wins, 8.4e9 against
    // 6.2e9; profiles/r02_sweep_z_3d_tree_larger_than_l2.txt)
    if (mode == kAuto && dt.blob_bytes > (size_t(48) << 20) && default_chunk(dt) < 32 * (int64_t)dt.view.n_leaves)
        return false;
    // 3D, measured crossover: 16 points per leaf (4096-leaf order-8 tree: 65 Ki points, 0.071 ms against 0.080 ms direct;
    // 2x ahead at 128 Ki), and from 16 Ki points on trees with few leaves (order 12: 0.038 against 0.069 ms)
    return block_bytes >= 1024 && n >= 16 * (int64_t)dt.view.n_leaves && n >= (1 << 14);
}

// Bucketed path: when the tree (nodes + coefficients) is too big for L2 the direct kernel's gathers go to DRAM,
// one or more sectors per level and per coefficient line; grouping the chunk's points by grid-cell bucket first
// turns them into cache hits.
bool use_bucketed(const Function *f, const DeviceTree &dt, int64_t n) {
    const int mode = f->mode.load();
    if (mode == kBucketed)
        return true;
    if (mode != kAuto)
        return false;
    return dt.blob_bytes > (size_t(48) << 20) && n >= (1 << 20);
}

int64_t default_bucket_chunk(const DeviceTree &dt) {
    if (const char *s = std::getenv("BAOBZI_B200_BCHUNK")) {
        long long v = std::atoll(s);
        if (v > 0)
            return v;
    }
    // The chunk's points and results (8 (dim + 1) bytes per point) should about fill the 126 MB L2: the evaluation's one
    // gather and one scatter per point then stay out of DRAM while the tree is streamed once per chunk by the staged kernel
    // (bucket_stage.cu).  Measured on the 2D 1M-leaf benchmark tree (profiles/r02_sweep_u_c4_chunk_stages.txt): 1.77e10
    // evals/s at 4 Mi points, 1.91e10 at 5 Mi, 1.94e10 at 5.5 Mi, 1.83e10 at 6 Mi, 1.69e10 at 7 Mi, 1.27e10 at 16 Mi; the
    // gathering kernel that serves trees the stage cannot hold is flat from 4 Mi to 16 Mi (1.31e10 - 1.32e10).
    return dt.view.dim <= 2 ? int64_t(11) << 19 : int64_t(1) << 22;
}

// evaluate n device-resident points on `stream`, chunking the binned path
int eval_device(Function *f, Replica *r, BinWorkspace &ws, const double *d_x, double *d_res, int64_t n,
                cudaStream_t stream) {
    int64_t launches = 0;
    int rc = 0;
    std::shared_ptr<StageProfile> prof_keep; // stays alive for this call even if baobzi_profile_end runs meanwhile
    {
        std::lock_guard<std::mutex> lock(f->prof_mu);
        prof_keep = f->profile;
    }
    StageProfile *prof = prof_keep.get();
    ws.prof = prof;
    const bool odd = misaligned_2d(r->dt, d_x);
    if (!odd && use_binned(f, r->dt, n)) {
        const int64_t fc = f->chunk.load();
        int64_t step = fc > 0 ? fc : default_chunk(r->dt);
        if (step > n)
            step = n;
        if (ws.chunk_capacity < step || ws.bucketed) {
            // (re)allocation: synchronous by nature (cudaMalloc / cudaFree), first call or growth only -- see baobzi.h
            cudaStreamSynchronize(stream);
            free_workspace(ws);
            rc = alloc_workspace(r->dt, step, ws);
        }
        ws.chunk_limit = step;
        ws.prof = prof; // (re)attach: a reallocation starts from a fresh workspace
        if (rc == 0)
            rc = launch_eval_binned(r->dt, ws, d_x, d_res, n, stream, &launches); // chunks of ws.chunk_capacity
    } else if (!odd && use_bucketed(f, r->dt, n)) {
        const int64_t fc = f->chunk.load();
        int64_t step = fc > 0 ? fc : default_bucket_chunk(r->dt);
        if (step > n)
            step = n;
        if (ws.chunk_capacity < step || !ws.bucketed) {
            cudaStreamSynchronize(stream);
            free_workspace(ws);
            rc = alloc_bucket_workspace(r->dt, step, ws);
        }
        ws.chunk_limit = step;
        ws.prof = prof;
        if (rc == 0)
            rc = launch_eval_bucketed(r->dt, ws, d_x, d_res, n, stream, &launches);
    } else {
        rc = launch_eval_direct(r->dt, d_x, d_res, nullptr, n, stream, &launches, prof);
    }
    f->launches += launches;
    return rc;
}

int env_int(const char *name, int dflt) {
    const char *s = std::getenv(name);
    return s && *s ? std::atoi(s) : dflt;
}

void fill_nan(double *res, int64_t n) {
    for (int64_t i = 0; i < n; ++i)
        res[i] = NAN;
}

int device_of_pointer(const void *p) { // -1: host (or unknown) memory
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return -1;
    }
    if (a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged)
        return a.device;
    return -1;
}

// ---- entry points stored in the handle (reference macros.h:29-51 generates 504 of these; here one set
// serves every (dim, order) because dispatch happens at kernel-launch time)
void fn_eval(const void *obj, const double *x, double *y) {
    Function *f = (Function *)obj;
    int dev = 0;
    cudaGetDevice(&dev);
    {
        std::lock_guard<std::mutex> lock(f->mu); // get_replica on another thread may grow the vector
        if (f->replicas.size() == 1)
            dev = f->replicas[0]->dt.device; // scalar calls stick to the replica that exists
    }
    Replica *r = get_replica(f, dev);
    const int od = f->host.output_dim < 1 ? 1 : (int)f->host.output_dim;
    ExecContext *c = r ? acquire(f, r) : nullptr;
    if (!c) {
        fill_nan(y, od);
        return;
    }
    DeviceGuard g(r->dt.device);
    bool bad = false;
    // scratch in mapped pinned memory: x[3] | completion word | y[od]
    if (c->pinned_doubles < (size_t)(4 + od)) {
        if (c->pinned)
            cudaFreeHost(c->pinned);
        c->pinned = nullptr;
        bad = report(cudaHostAlloc((void **)&c->pinned, sizeof(double) * (4 + od), cudaHostAllocMapped), "pinned alloc");
        c->pinned_doubles = bad ? 0 : (size_t)(4 + od);
        c->scalar_seq = 0;
        if (!bad)
            std::memset(c->pinned, 0, sizeof(double) * (4 + od));
    }
    if (!bad) {
        // zero-copy: the kernel reads the point from and writes the values to mapped pinned memory, then publishes a
        // sequence number there; the host polls that word (no driver call on the way back) and only falls back to a
        // stream synchronisation when it does not show up -- which is also how a launch error surfaces
        static const bool spin = env_int("BAOBZI_B200_SCALAR_SPIN", 1) != 0;
        volatile unsigned long long *done = reinterpret_cast<volatile unsigned long long *>(c->pinned + 3);
        double *yv = c->pinned + 4;
        for (int j = 0; j < f->host.dim; ++j)
            c->pinned[j] = x[j];
        for (int k = 0; k < od; ++k)
            yv[k] = NAN;
        const unsigned long long seq = ++c->scalar_seq;
        int64_t launches = 0;
        int rc = launch_eval_scalar(r->dt, c->pinned, yv, c->stream, &launches,
                                    spin ? const_cast<unsigned long long *>(done) : nullptr, seq);
        f->launches += launches;
        bad = report((cudaError_t)rc, "scalar eval launch");
        bool seen = false;
        if (!bad && spin) {
            for (int it = 0; it < 2000000 && !seen; ++it) // ~ a few hundred ms at most, normally < 10 us
                seen = *done == seq;
            std::atomic_thread_fence(std::memory_order_acquire);
        }
        if (!bad && !seen)
            bad = report(cudaStreamSynchronize(c->stream), "scalar eval");
        if (!bad)
            for (int k = 0; k < od; ++k)
                y[k] = yv[k];
    }
    if (bad)
        fill_nan(y, od);
    release(f, c);
}

// ---- host-pointer path ---------------------------------------------------------------------------------------
// A call with host pointers is cut into chunks that a few WORKER threads take from a shared counter.  Each worker owns
// one execution context (stream, device chunk buffers, sort workspace) on one GPU and runs, per chunk,
//     [memcpy caller -> pinned staging]  H2D  kernels  D2H  [memcpy pinned staging -> caller]
// The bracketed steps only exist for PAGEABLE caller memory -- what the reference's callers hand over (malloc in
// examples/C/baobzi_timing.c:44-49, numpy in src/python/__init__.py:126-131): cudaMemcpyAsync from pageable memory is
// staged by the driver through one small buffer and is synchronous with the host, so the pipeline would degenerate
// to one copy at a time at a fraction of the link rate; one host thread moves ~10 GB/s, the link needs ~55, hence
// several workers.  With pinned (or cudaHostRegister'ed) memory three workers per GPU keep both directions of the
// link and the kernels busy.  Workers on several GPUs (baobzi_set_host_devices / BAOBZI_B200_DEVICES) share the same
// counter, so the chunks balance themselves over the devices; the tree is replicated lazily, there is no collective.
struct HostJob {
    Function *f;
    const double *x;
    double *res;
    int64_t n, chunk, n_chunks;
    int dim;
    bool stage_x, stage_res;
    std::atomic<int64_t> next{0};
    std::atomic<bool> bad{false};
};

bool ensure_chunk_buffers(ExecContext *c, const DeviceTree &dt, int64_t points, bool stage_x, bool stage_res) {
    if (c->slot_points < points) {
        bz_free(c->d_x);
        bz_free(c->d_res);
        c->d_x = c->d_res = nullptr;
        c->slot_points = 0;
        if (report(bz_malloc((void **)&c->d_x, sizeof(double) * points * dt.view.dim), "chunk alloc") ||
            report(bz_malloc((void **)&c->d_res, sizeof(double) * points), "chunk alloc"))
            return false;
        c->slot_points = points;
    }
    if ((stage_x || stage_res) && c->staged_points < points) {
        if (c->h_x)
            cudaFreeHost(c->h_x);
        if (c->h_res)
            cudaFreeHost(c->h_res);
        c->h_x = c->h_res = nullptr;
        c->staged_points = 0;
        if (report(cudaHostAlloc((void **)&c->h_x, sizeof(double) * points * dt.view.dim, cudaHostAllocDefault), "pinned staging alloc") ||
            report(cudaHostAlloc((void **)&c->h_res, sizeof(double) * points, cudaHostAllocDefault), "pinned staging alloc"))
            return false;
        c->staged_points = points;
    }
    return true;
}

void host_worker(HostJob *job, int dev) {
    Function *f = job->f;
    if (report(cudaSetDevice(dev), "cudaSetDevice")) {
        job->bad = true;
        return;
    }
    Replica *r = get_replica(f, dev);
    ExecContext *c = r ? acquire(f, r) : nullptr;
    if (!c || !ensure_chunk_buffers(c, r->dt, job->chunk, job->stage_x, job->stage_res)) {
        job->bad = true;
        if (c)
            release(f, c);
        return;
    }
    const int dim = job->dim;
    cudaStream_t st = c->stream;
    for (;;) {
        const int64_t k = job->next.fetch_add(1);
        if (k >= job->n_chunks || job->bad.load())
            break;
        const int64_t o = k * job->chunk, m = std::min(job->chunk, job->n - o);
        const double *src = job->x + o * dim;
        if (job->stage_x) {
            std::memcpy(c->h_x, src, sizeof(double) * m * dim);
            src = c->h_x;
        }
        double *dst = job->stage_res ? c->h_res : job->res + o;
        bool bad = report(cudaMemcpyAsync(c->d_x, src, sizeof(double) * m * dim, cudaMemcpyHostToDevice, st), "H2D copy");
        if (!bad)
            bad = report((cudaError_t)eval_device(f, r, c->ws, c->d_x, c->d_res, m, st), "eval launch");
        if (!bad)
            bad = report(cudaMemcpyAsync(dst, c->d_res, sizeof(double) * m, cudaMemcpyDeviceToHost, st), "D2H copy");
        bad = report(cudaStreamSynchronize(st), "eval_multi") || bad;
        if (bad) {
            job->bad = true;
            break;
        }
        if (job->stage_res)
            std::memcpy(job->res + o, c->h_res, sizeof(double) * m);
    }
    release(f, c);
}

bool is_pageable(const void *p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return true;
    }
    return a.type == cudaMemoryTypeUnregistered;
}

void fn_eval_multi(const void *obj, const double *x, double *res, int ntrg) {
    Function *f = (Function *)obj;
    if (ntrg <= 0)
        return;
    // Device pointers handed to the classic entry point are evaluated in place (zero copy).
    const int xdev = device_of_pointer(x), rdev = device_of_pointer(res);
    if (xdev >= 0 && rdev == xdev) {
        Replica *r = get_replica(f, xdev);
        ExecContext *c = r ? acquire(f, r) : nullptr;
        if (!c)
            return;
        DeviceGuard g(xdev);
        int rc = eval_device(f, r, c->ws, x, res, ntrg, c->stream);
        report((cudaError_t)rc, "eval_multi (device pointers)");
        report(cudaStreamSynchronize(c->stream), "eval_multi (device pointers)");
        release(f, c);
        return;
    }

    int dev0 = 0;
    cudaGetDevice(&dev0);
    const int64_t n = ntrg;
    const int dim = f->host.dim;
    HostJob job;
    job.f = f;
    job.x = x;
    job.res = res;
    job.n = n;
    job.dim = dim;
    // pageable memory is staged through pinned buffers once the call is big enough for that to pay (below, the
    // driver's own staging of a single cudaMemcpyAsync is as good); BAOBZI_B200_STAGE=0 / 1 forces it off / on
    const int stage_env = env_int("BAOBZI_B200_STAGE", -1);
    const bool big = n * (dim + 1) * (int64_t)sizeof(double) >= (int64_t(8) << 20);
    job.stage_x = stage_env == 0 ? false : (is_pageable(x) && (big || stage_env == 1));
    job.stage_res = stage_env == 0 ? false : (is_pageable(res) && (big || stage_env == 1));
    // alternative for pageable memory (BAOBZI_B200_HOSTREG=1): pin the caller's pages for the duration of the call.
    // Pinning costs about as much as copying (the kernel touches every page) and is not overlapped with anything, so
    // it only wins when the same buffers come back call after call; measured in INTEGRATION.md, off by default.
    bool reg_x = false, reg_res = false;
    if (env_int("BAOBZI_B200_HOSTREG", 0) != 0 && big) {
        if (job.stage_x && cudaHostRegister(const_cast<double *>(x), sizeof(double) * n * dim, cudaHostRegisterPortable) == cudaSuccess)
            reg_x = true, job.stage_x = false;
        if (job.stage_res && cudaHostRegister(res, sizeof(double) * n, cudaHostRegisterPortable) == cudaSuccess)
            reg_res = true, job.stage_res = false;
        cudaGetLastError();
    }
    // devices: the current one first, then the others in order
    int visible = 1;
    cudaGetDeviceCount(&visible);
    int n_dev = f->host_devices <= 0 ? visible : std::min(f->host_devices, visible);
    if (n < (int64_t(1) << 22))
        n_dev = 1; // not worth waking other GPUs (and replicating the tree) for
    // workers per device: 3 keep H2D, kernels and D2H of a pinned-memory call overlapped; staging needs more host
    // threads than that to feed the link (measured on B200 hosts)
    int per_dev = (job.stage_x || job.stage_res) ? 8 : 3; // measured plateau (profiles/r02_host_path_ab.txt: 2 / 4 / 8 / 12 / 16 workers)
    per_dev = env_int("BAOBZI_B200_HOST_WORKERS", per_dev);
    per_dev = std::max(1, std::min(per_dev, 16));
    const int hw = (int)std::thread::hardware_concurrency();
    if (hw > 0 && n_dev * per_dev > std::max(4, hw))
        per_dev = std::max(1, std::max(4, hw) / n_dev);
    // staging is bound by the host's memory system, not by the number of links: beyond ~12 copying threads in total the
    // throughput FALLS (2 GPUs x 8 workers: 1.05e9 evals/s against 1.63e9 for 1 x 8, profiles/r02_host_path_ab_2gpu.txt)
    if ((job.stage_x || job.stage_res) && n_dev * per_dev > 12 && !std::getenv("BAOBZI_B200_HOST_WORKERS"))
        per_dev = std::max(1, 12 / n_dev);
    // chunk: 4 Mi points, or less so that every worker gets something
    int64_t chunk = int64_t(1) << 22;
    const int64_t workers_wanted = (int64_t)n_dev * per_dev;
    if (n < chunk * workers_wanted)
        chunk = std::max<int64_t>((n + workers_wanted - 1) / workers_wanted, std::min<int64_t>(n, 1 << 16));
    job.chunk = chunk;
    job.n_chunks = (n + chunk - 1) / chunk;
    const int n_workers = (int)std::min<int64_t>(workers_wanted, job.n_chunks);
    if (n_workers <= 1) {
        host_worker(&job, dev0);
        cudaSetDevice(dev0);
    } else {
        std::vector<std::thread> th;
        th.reserve(n_workers);
        for (int w = 0; w < n_workers; ++w)
            th.emplace_back(host_worker, &job, (dev0 + w % n_dev) % visible);
        for (auto &t : th)
            t.join();
    }
    if (reg_x)
        cudaHostUnregister(const_cast<double *>(x));
    if (reg_res)
        cudaHostUnregister(res);
    if (job.bad.load())
        fill_nan(res, n);
}

void fn_save(const void *obj, const char *filename) {
    const Function *f = (const Function *)obj;
    std::string err;
    if (!write_baobzi_file(filename, f->host, err))
        std::fprintf(stderr, "Baobzi save error: %s\n", err.c_str());
}

void fn_stats(void *obj) {
    Function *f = (Function *)obj;
    const HostTree &h = f->host;
    // same lines as the reference's print_stats (template.hpp:582-603), plus the device footprint
    size_t host_bytes = h.coeffs.size() * 8 + (size_t)h.n_nodes() * (16 * h.dim + 16) + (size_t)h.n_bins() * 4;
    std::printf("Baobzi function mapping %d to %u\n", h.dim, h.output_dim);
    std::printf("Tree represented by %lld nodes, of which %lld are leaves\n", (long long)h.n_nodes(),
                (long long)h.n_leaves());
    std::printf("Nodes are distributed across %lld subtrees at an initial depth of %d with a maximum subtree depth of %d\n",
                (long long)h.n_bins(), h.base_depth, h.max_subtree_depth);
    std::printf("Total function evaluations required for fit: %llu\n", (unsigned long long)h.n_func_evals);
    std::printf("Total time to create tree: %u milliseconds\n", h.fit_ms);
    std::printf("Approximate memory usage of tree: %g MiB\n", (double)host_bytes / (1024.0 * 1024.0));
    std::lock_guard<std::mutex> lock(f->mu);
    for (auto &r : f->replicas)
        std::printf("Device replica on GPU %d: %g MiB (64-byte node records + coefficients)\n", r->dt.device,
                    (double)r->dt.blob_bytes / (1024.0 * 1024.0));
    std::fflush(stdout);
}

void fn_free(void *obj) {
    Function *f = (Function *)obj;
    for (auto &r : f->replicas) {
        DeviceGuard g(r->dt.device);
        for (auto &c : r->contexts)
            destroy_context(*c);
        if (r->shard_stream)
            cudaStreamDestroy(r->shard_stream);
        free_workspace(r->shard_ws);
        if (r->ws_done)
            cudaEventDestroy(r->ws_done);
        if (r->dot_partial)
            cudaFree(r->dot_partial);
        if (r->dot_tmp)
            cudaFree(r->dot_tmp);
        free_tree(r->dt);
    }
    delete f;
}

baobzi_t make_handle(std::unique_ptr<Function> f) {
    int dev = 0;
    if (report(cudaGetDevice(&dev), "device query (is a CUDA device visible? there is no CPU evaluation path)"))
        return nullptr;
    f->mode.store(env_mode());
    if (const char *g = std::getenv("BAOBZI_B200_L2_FETCH")) // experiment knob: L2 fetch granularity hint (32 / 64 / 128 bytes)
        if (std::atoi(g) > 0)
            cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, (size_t)std::atoi(g));
    if (const char *d = std::getenv("BAOBZI_B200_DEVICES"))
        f->host_devices = !std::strcmp(d, "all") ? 0 : std::max(1, std::atoi(d));
    if (!get_replica(f.get(), dev))
        return nullptr;
    baobzi_t res = (baobzi_t)std::malloc(sizeof(baobzi_struct));
    res->DIM = f->host.dim;
    res->ORDER = f->host.order;
    res->OUTPUT_DIM = (int)f->host.output_dim;
    res->eval = &fn_eval;
    res->eval_multi = &fn_eval_multi;
    res->save = &fn_save;
    res->stats = &fn_stats;
    res->free = &fn_free;
    res->obj = f.release();
    return res;
}

} // namespace

// noalloc.cu (baobzi_eval_multi_device_noalloc): the tree descriptor of the replica on `dev` if there is one.  Never
// creates a replica, so it never allocates.
bool find_replica_view(const void *obj, int dev, TreeView *out) {
    Function *f = (Function *)obj;
    std::lock_guard<std::mutex> lock(f->mu);
    for (auto &r : f->replicas)
        if (r->dt.device == dev) {
            *out = r->dt.view;
            return true;
        }
    return false;
}
} // namespace bz

// ================================================================================================
// exported C symbols
// ================================================================================================
using bz::Function;

extern "C" {

const baobzi_input_t baobzi_input_default; // zero + C++ member defaults, as reference baobzi.cpp:16

const char *baobzi_b200_version(void) { return "baobzi-b200 0.1 (sm_100a)"; }

void baobzi_eval(const baobzi_t func, const double *x, double *y) { func->eval(func->obj, x, y); }

void baobzi_eval_multi(const baobzi_t func, const double *x, double *res, int ntrg) {
    func->eval_multi(func->obj, x, res, ntrg);
}

void baobzi_save(const baobzi_t func, const char *filename) { func->save(func->obj, filename); }

// extern "C" entry points must not let a C++ exception (bad_alloc on a huge tree, ...) unwind into a C caller
static baobzi_t restore_impl(const char *filename);
static baobzi_t init_impl(const baobzi_input_t *input, const double *center, const double *half_length);

baobzi_t baobzi_restore(const char *filename) {
    try {
        return restore_impl(filename);
    } catch (const std::exception &e) {
        std::fprintf(stderr, "Baobzi restore error: Unable to restore from '%s' (%s)\n", filename, e.what());
    } catch (...) {
        std::fprintf(stderr, "Baobzi restore error: Unable to restore from '%s'\n", filename);
    }
    return nullptr;
}

baobzi_t baobzi_init(const baobzi_input_t *input, const double *center, const double *half_length) {
    try {
        return init_impl(input, center, half_length);
    } catch (const std::exception &e) { // reference baobzi.cpp:178-182 prints the exception text and returns NULL
        std::fprintf(stderr, "%s\n", e.what());
    } catch (...) {
        std::fprintf(stderr, "Baobzi error: unknown exception during initialization\n");
    }
    return nullptr;
}

static baobzi_t restore_impl(const char *filename) {
    auto f = std::make_unique<Function>();
    std::string err;
    if (!bz::read_baobzi_file(filename, f->host, err)) {
        // reference baobzi.cpp:94-114: silent NULL on a version mismatch, message otherwise
        if (err.find("version") == std::string::npos) {
            if (err.find("(DIM, ORDER)") != std::string::npos)
                std::fprintf(stderr, "BAOBZI ERROR: %s\n", err.c_str());
            else
                std::fprintf(stderr, "Baobzi restore error: Unable to restore from '%s' (%s)\n", filename, err.c_str());
        }
        return nullptr;
    }
    return bz::make_handle(std::move(f));
}

static baobzi_t init_impl(const baobzi_input_t *input, const double *center, const double *half_length) {
    // argument checks in the order of reference baobzi.cpp:148-157
    if (input->tol <= 0.0) {
        std::fprintf(stderr, "Baobzi error: Unable to initialize Baobzi due to invalid 'tol' parameter. Please supply "
                             "something greater than zero.\n");
        return nullptr;
    }
    if (!input->func) {
        std::fprintf(stderr, "BAOBZI ERROR: Unable to initialize Baobzi due to empty or invalid 'func' parameter. "
                             "Please supply a valid function to fit.\n");
        return nullptr;
    }
    if (!bz::supported_dim_order(input->dim, input->order)) { // reference baobzi.cpp:170-176
        std::fprintf(stderr, "Baobzi error: Unable to initialize Baobzi function with variables (DIM, ORDER): (%d, %d)\n",
                     input->dim, input->order);
        return nullptr;
    }
    auto f = std::make_unique<Function>();
    std::string err;
    if (!bz::fit_tree(*input, center, half_length, f->host, err)) { // reference baobzi.cpp:178-182
        std::fprintf(stderr, "%s\n", err.c_str());
        return nullptr;
    }
    return bz::make_handle(std::move(f));
}

void baobzi_stats(baobzi_t func) {
    if (!func)
        return;
    func->stats(func->obj);
}

baobzi_t baobzi_free(baobzi_t func) {
    if (!func)
        return nullptr;
    if (func->obj)
        func->free(func->obj);
    std::free(func);
    return nullptr;
}

// ---------------------------------------------------------------------------- device extension
int baobzi_eval_multi_device(const baobzi_t func, const double *d_x, double *d_res, int64_t ntrg, void *cuda_stream) {
    Function *f = (Function *)func->obj;
    if (ntrg <= 0)
        return 0;
    int dev = bz::device_of_pointer(d_x);
    if (dev < 0)
        return (int)cudaErrorInvalidDevicePointer;
    bz::Replica *r = bz::get_replica(f, dev);
    if (!r)
        return (int)cudaErrorMemoryAllocation;
    bz::DeviceGuard g(dev);
    // the binned workspace of the per-device shard slot is reused; calls on one handle+device are
    // serialised on the caller's stream by construction (one workspace), so guard it
    std::lock_guard<std::mutex> lock(f->mu);
    bz::workspace_begin(r, (cudaStream_t)cuda_stream);
    const int rc = bz::eval_device(f, r, r->shard_ws, d_x, d_res, ntrg, (cudaStream_t)cuda_stream);
    bz::workspace_end(r, (cudaStream_t)cuda_stream);
    return rc;
}

int baobzi_eval_multi_device_all(const baobzi_t func, const double *d_x, double *d_res, int64_t ntrg, void *cuda_stream) {
    Function *f = (Function *)func->obj;
    if (ntrg <= 0)
        return 0;
    if (func->DIM != 1)
        return (int)cudaErrorNotSupported; // the reference's multi-output fit is only meaningful in 1D
    int dev = bz::device_of_pointer(d_x);
    if (dev < 0)
        return (int)cudaErrorInvalidDevicePointer;
    bz::Replica *r = bz::get_replica(f, dev);
    if (!r)
        return (int)cudaErrorMemoryAllocation;
    bz::DeviceGuard g(dev);
    int64_t launches = 0;
    const int rc = bz::launch_eval_all_1d(r->dt, d_x, d_res, ntrg, (cudaStream_t)cuda_stream, &launches);
    f->launches += launches;
    return rc;
}

// grow a device buffer of doubles (the old one may still be in use by work queued on `stream`)
static int reserve_doubles(double *&buf, int64_t &cap, int64_t want, cudaStream_t stream) {
    if (cap >= want)
        return 0;
    if (buf) {
        cudaStreamSynchronize(stream);
        cudaFree(buf);
        buf = nullptr;
        cap = 0;
    }
    cudaError_t e = cudaMalloc(&buf, (size_t)want * sizeof(double));
    if (e != cudaSuccess)
        return (int)e;
    cap = want;
    return 0;
}

int baobzi_eval_multi_device_dot(const baobzi_t func, const double *d_x, const double *d_w, int64_t ntrg, double *d_out,
                                 void *cuda_stream) {
    Function *f = (Function *)func->obj;
    cudaStream_t stream = (cudaStream_t)cuda_stream;
    if (ntrg <= 0)
        return (int)cudaMemsetAsync(d_out, 0, sizeof(double), stream); // empty sum
    int dev = bz::device_of_pointer(d_x);
    if (dev < 0)
        return (int)cudaErrorInvalidDevicePointer;
    bz::Replica *r = bz::get_replica(f, dev);
    if (!r)
        return (int)cudaErrorMemoryAllocation;
    bz::DeviceGuard g(dev);
    std::lock_guard<std::mutex> lock(f->mu);
    const int G = bz::dot_grid(r->dt);
    int64_t launches = 0;
    int rc = 0;
    bz::workspace_begin(r, stream);
    const bool odd = bz::misaligned_2d(r->dt, d_x);
    const bool binned = !odd && bz::use_binned(f, r->dt, ntrg), bucketed = !odd && !binned && bz::use_bucketed(f, r->dt, ntrg);
    if (!binned && !bucketed) { // fused: the values never leave registers
        rc = reserve_doubles(r->dot_partial, r->dot_partial_cap, G, stream);
        if (rc == 0)
            rc = bz::launch_eval_dot(r->dt, d_x, d_w, ntrg, r->dot_partial, stream, &launches);
        if (rc == 0)
            rc = bz::launch_sum_partials(r->dot_partial, G, d_out, stream, &launches);
    } else { // grouped evaluation chunk by chunk into an L2-sized scratch array, consumed at once
        const int64_t fc = f->chunk.load();
        int64_t step = fc > 0 ? fc : (binned ? bz::default_chunk(r->dt) : bz::default_bucket_chunk(r->dt));
        if (step > ntrg)
            step = ntrg;
        const int64_t n_chunks = (ntrg + step - 1) / step;
        rc = reserve_doubles(r->dot_partial, r->dot_partial_cap, n_chunks * G, stream);
        if (rc == 0)
            rc = reserve_doubles(r->dot_tmp, r->dot_tmp_cap, step, stream);
        for (int64_t c = 0; c < n_chunks && rc == 0; ++c) {
            const int64_t off = c * step, m = std::min(step, ntrg - off);
            rc = bz::eval_device(f, r, r->shard_ws, d_x + off * func->DIM, r->dot_tmp, m, stream);
            if (rc == 0)
                rc = bz::launch_dot(r->dt, r->dot_tmp, d_w ? d_w + off : nullptr, m, r->dot_partial + c * G, stream,
                                    &launches);
        }
        if (rc == 0)
            rc = bz::launch_sum_partials(r->dot_partial, n_chunks * G, d_out, stream, &launches);
    }
    bz::workspace_end(r, stream);
    f->launches += launches;
    return rc;
}

int baobzi_device_view(const baobzi_t func, int device, void *view, size_t view_bytes) {
    Function *f = (Function *)func->obj;
    if (!view || view_bytes != sizeof(bz::TreeView))
        return (int)cudaErrorInvalidValue; // header / library mismatch
    bz::Replica *r = bz::get_replica(f, device);
    if (!r)
        return (int)cudaErrorMemoryAllocation;
    std::memcpy(view, &r->dt.view, sizeof(bz::TreeView));
    return 0;
}

int baobzi_node_index_multi_device(const baobzi_t func, const double *d_x, int64_t *d_idx, int64_t ntrg,
                                   void *cuda_stream) {
    Function *f = (Function *)func->obj;
    if (ntrg <= 0)
        return 0;
    int dev = bz::device_of_pointer(d_x);
    if (dev < 0)
        return (int)cudaErrorInvalidDevicePointer;
    bz::Replica *r = bz::get_replica(f, dev);
    if (!r)
        return (int)cudaErrorMemoryAllocation;
    bz::DeviceGuard g(dev);
    int64_t launches = 0;
    int rc = bz::launch_eval_direct(r->dt, d_x, nullptr, d_idx, ntrg, (cudaStream_t)cuda_stream, &launches);
    f->launches += launches;
    return rc;
}

int baobzi_node_index_multi(const baobzi_t func, const double *x, int64_t *idx, int64_t ntrg) {
    Function *f = (Function *)func->obj;
    if (ntrg <= 0)
        return 0;
    int dev = 0;
    cudaGetDevice(&dev);
    bz::Replica *r = bz::get_replica(f, dev);
    if (!r)
        return (int)cudaErrorMemoryAllocation;
    double *d_x = nullptr;
    int64_t *d_idx = nullptr;
    cudaError_t e = cudaMalloc((void **)&d_x, sizeof(double) * ntrg * f->host.dim);
    if (e == cudaSuccess)
        e = cudaMalloc((void **)&d_idx, sizeof(int64_t) * ntrg);
    if (e == cudaSuccess)
        e = cudaMemcpy(d_x, x, sizeof(double) * ntrg * f->host.dim, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) {
        int64_t launches = 0;
        e = (cudaError_t)bz::launch_eval_direct(r->dt, d_x, nullptr, d_idx, ntrg, nullptr, &launches);
        f->launches += launches;
    }
    if (e == cudaSuccess)
        e = cudaMemcpy(idx, d_idx, sizeof(int64_t) * ntrg, cudaMemcpyDeviceToHost);
    cudaFree(d_x);
    cudaFree(d_idx);
    return (int)e;
}

int baobzi_eval_multi_sharded(const baobzi_t func, int n_dev, const int *devices, const double *const *d_x,
                              double *const *d_res, const int64_t *ntrg) {
    Function *f = (Function *)func->obj;
    int rc = 0;
    for (int k = 0; k < n_dev && rc == 0; ++k) {
        bz::Replica *r = bz::get_replica(f, devices[k]);
        if (!r)
            return (int)cudaErrorMemoryAllocation;
        bz::DeviceGuard g(devices[k]);
        if (!r->shard_stream)
            rc = (int)cudaStreamCreateWithFlags(&r->shard_stream, cudaStreamNonBlocking);
        if (rc == 0 && ntrg[k] > 0) {
            std::lock_guard<std::mutex> lock(f->mu);
            bz::workspace_begin(r, r->shard_stream);
            rc = bz::eval_device(f, r, r->shard_ws, d_x[k], d_res[k], ntrg[k], r->shard_stream);
            bz::workspace_end(r, r->shard_stream);
        }
    }
    return rc;
}

int baobzi_sync(const baobzi_t func) {
    Function *f = (Function *)func->obj;
    int rc = 0;
    std::vector<std::pair<int, cudaStream_t>> streams; // snapshot under the lock, wait outside it
    {
        std::lock_guard<std::mutex> lock(f->mu);
        for (auto &r : f->replicas)
            if (r->shard_stream)
                streams.emplace_back(r->dt.device, r->shard_stream);
    }
    for (auto &ds : streams) {
        bz::DeviceGuard g(ds.first);
        int e = (int)cudaStreamSynchronize(ds.second);
        rc = rc ? rc : e;
    }
    return rc;
}

int baobzi_set_mode(baobzi_t func, int mode, int64_t chunk) {
    Function *f = (Function *)func->obj;
    if (mode < 0 || mode > 3)
        return 1;
    f->mode.store(mode);
    f->chunk.store(chunk > 0 ? chunk : 0);
    return 0;
}

int baobzi_tree_info(const baobzi_t func, int64_t info[5]) {
    Function *f = (Function *)func->obj;
    info[0] = f->host.n_nodes();
    info[1] = f->host.n_leaves();
    info[2] = f->host.n_bins();
    info[3] = (int64_t)f->host.coeffs.size();
    std::lock_guard<std::mutex> lock(f->mu);
    info[4] = f->replicas.empty() ? 0 : (int64_t)f->replicas[0]->dt.blob_bytes;
    return 0;
}

int baobzi_set_host_devices(baobzi_t func, int n_devices) {
    Function *f = (Function *)func->obj;
    if (n_devices < 0)
        return 1;
    f->host_devices = n_devices; // 0 = every visible device
    return 0;
}

int baobzi_b200_guard_check(int64_t *bad_bytes, int64_t *buffers) {
    return bz::guard_check(bad_bytes, buffers);
}

int64_t baobzi_launch_count(const baobzi_t func) { return ((Function *)func->obj)->launches.load(); }

int baobzi_profile_begin(baobzi_t func) {
    Function *f = (Function *)func->obj;
    std::lock_guard<std::mutex> lock(f->prof_mu);
    f->profile = std::make_shared<bz::StageProfile>();
    return 0;
}

int baobzi_profile_end(baobzi_t func, double *ms, int64_t *launches, int64_t *points) {
    Function *f = (Function *)func->obj;
    cudaError_t e = cudaDeviceSynchronize();
    std::lock_guard<std::mutex> lock(f->prof_mu);
    if (!f->profile)
        return (int)cudaErrorInvalidValue;
    f->profile->collect(ms, launches, points);
    f->profile.reset();
    return (int)e;
}

int baobzi_b200_fp64_peak(double *tflops) { return bz::measure_fp64_peak(tflops, 5); }
int baobzi_b200_fp64_peak_sustained(double *tflops, double seconds) { return bz::measure_fp64_peak_sustained(tflops, seconds); }

} // extern "C"
