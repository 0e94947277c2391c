// Native per-GPU runner of the frame-sharded front end (one worker per GPU, frames b -> GPU
// b mod G, inputs staged, no NCCL).  The reference pins everything to cuda:0 (train.py:27, 190, 276).
//
// A runner owns ONE host thread and a ring of slots; a slot is what the host language prepared once: device + pinned staging
// buffers for a batch of frames, a stream, and an instantiated CUDA graph of the step (offsets H2D, voxelise, decorate / PFN,
// canvas fill, counts D2H).  pf_runner_submit only enqueues pointers and returns a ticket; the runner's thread uploads the
// frames (pf_upload_frames: pageable frames through the pinned staging buffer), launches the slot's graph and records an
// event; pf_runner_wait blocks on that event.  Nothing here needs the interpreter, so eight GPUs are driven by eight native
// threads and the Python side costs one call per GPU and batch (a Python worker thread per GPU serialises on the
// interpreter lock: 4.2 G points/s at 8 GPUs against 26 G on the device, profiles/r2_sharded_pinned_n8.json).
#include <condition_variable>
#include <deque>
#include <mutex>
#include <new>
#include <string>
#include <thread>
#include <vector>

#include "pf_internal.cuh"

namespace {

struct Slot {
    cudaGraphExec_t exec = nullptr;
    cudaStream_t stream = nullptr;
    void *d_pts = nullptr, *h_stage = nullptr;
    int32_t *h_offs = nullptr;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    bool ready = false, used = false;
};

struct Job {
    long long ticket;
    std::vector<const void *> ptrs;
    std::vector<int64_t> rows;
    std::vector<int32_t> kinds;
};

}  // namespace

struct pf_runner {
    int device = 0, frames_per_batch = 0, row_bytes = 0, stage_threads = 1;
    std::vector<Slot> slots;
    std::thread th;
    std::mutex m;
    std::condition_variable cv_job, cv_done;
    std::deque<Job> q;
    long long next_ticket = 0, launched = -1;
    bool stop = false;
    struct Failure {
        long long ticket;
        int status;          // pf_status or cudaError_t
        std::string what;
    };
    std::deque<Failure> failures;   // batches the worker thread could not launch (kept until waited for)
};

namespace {

thread_local std::string g_runner_err;

void worker(pf_runner *r)
{
    cudaSetDevice(r->device);
    for (;;) {
        Job job;
        {
            std::unique_lock<std::mutex> lk(r->m);
            r->cv_job.wait(lk, [&] { return r->stop || !r->q.empty(); });
            if (r->q.empty()) return;      // stop requested and nothing left
            job = std::move(r->q.front());
            r->q.pop_front();
        }
        Slot &s = r->slots[(size_t)(job.ticket % (long long)r->slots.size())];
        int rc = 0;
        std::string what;
        if (!s.ready) {
            rc = PF_ERR_STATE, what = "pf_runner: slot was never set";
        } else {
            if (s.used) cudaEventSynchronize(s.e1);      // the batch that last ran in this slot must have left the GPU
            cudaError_t e = cudaEventRecord(s.e0, s.stream);
            if (e == cudaSuccess) {
                rc = pf_upload_frames(s.d_pts, s.h_stage, job.ptrs.data(), job.rows.data(), job.kinds.data(), (int)job.ptrs.size(),
                                      r->frames_per_batch, r->row_bytes, r->stage_threads, s.h_offs, s.stream);
                if (rc != 0) what = pf_last_error_string();
            }
            if (e == cudaSuccess && rc == 0) e = cudaGraphLaunch(s.exec, s.stream);
            if (e == cudaSuccess && rc == 0) e = cudaEventRecord(s.e1, s.stream);
            if (e != cudaSuccess) rc = (int)e, what = std::string("pf_runner: ") + cudaGetErrorName(e);
            s.used = true;
        }
        {
            std::lock_guard<std::mutex> lk(r->m);
            if (rc != 0) r->failures.push_back({job.ticket, rc, what});
            r->launched = job.ticket;
        }
        r->cv_done.notify_all();
    }
}

}  // namespace

extern "C" {

const char *pf_runner_last_error(void) { return g_runner_err.c_str(); }

int pf_runner_create(int device, int n_slots, int frames_per_batch, int row_bytes, int stage_threads, pf_runner **out)
{
    if (!out || n_slots < 1 || n_slots > 64 || frames_per_batch < 1 || row_bytes < 4) {
        g_runner_err = "pf_runner_create: bad argument";
        return PF_ERR_INVALID_ARGUMENT;
    }
    pf_runner *r = new (std::nothrow) pf_runner();
    if (!r) return PF_ERR_OUT_OF_MEMORY;
    r->device = device, r->frames_per_batch = frames_per_batch, r->row_bytes = row_bytes;
    r->stage_threads = stage_threads < 1 ? 1 : stage_threads;
    r->slots.resize((size_t)n_slots);
    int prev = -1;
    cudaGetDevice(&prev);
    cudaError_t e = cudaSetDevice(device);
    for (auto &s : r->slots) {
        if (e == cudaSuccess) e = cudaEventCreate(&s.e0);
        if (e == cudaSuccess) e = cudaEventCreate(&s.e1);
    }
    if (prev >= 0) cudaSetDevice(prev);
    if (e != cudaSuccess) {
        g_runner_err = std::string("pf_runner_create: ") + cudaGetErrorName(e);
        delete r;
        return (int)e;
    }
    r->th = std::thread(worker, r);
    *out = r;
    return PF_OK;
}

int pf_runner_set_slot(pf_runner *r, int slot, void *graph_exec, void *stream, void *device_points, void *pinned_stage,
                       int32_t *pinned_offsets)
{
    if (!r || slot < 0 || slot >= (int)r->slots.size() || !graph_exec || !device_points || !pinned_offsets) {
        g_runner_err = "pf_runner_set_slot: bad argument";
        return PF_ERR_INVALID_ARGUMENT;
    }
    std::lock_guard<std::mutex> lk(r->m);
    Slot &s = r->slots[(size_t)slot];
    s.exec = static_cast<cudaGraphExec_t>(graph_exec), s.stream = static_cast<cudaStream_t>(stream);
    s.d_pts = device_points, s.h_stage = pinned_stage, s.h_offs = pinned_offsets;
    s.ready = true;
    return PF_OK;
}

long long pf_runner_submit(pf_runner *r, const void *const *frames, const int64_t *rows, const int32_t *kinds, int n_frames)
{
    if (!r || n_frames < 0 || n_frames > r->frames_per_batch || (n_frames > 0 && (!frames || !rows || !kinds))) {
        g_runner_err = "pf_runner_submit: bad argument";
        return PF_ERR_INVALID_ARGUMENT;
    }
    Job job;
    job.ptrs.assign(frames, frames + n_frames);
    job.rows.assign(rows, rows + n_frames);
    job.kinds.assign(kinds, kinds + n_frames);
    long long t;
    {
        std::lock_guard<std::mutex> lk(r->m);
        if (r->stop) {
            g_runner_err = "pf_runner_submit: runner is shutting down";
            return PF_ERR_STATE;
        }
        t = job.ticket = r->next_ticket++;
        r->q.push_back(std::move(job));
    }
    r->cv_job.notify_one();
    return t;
}

int pf_runner_wait(pf_runner *r, long long ticket, float *device_ms)
{
    if (!r || ticket < 0) {
        g_runner_err = "pf_runner_wait: bad argument";
        return PF_ERR_INVALID_ARGUMENT;
    }
    {
        std::unique_lock<std::mutex> lk(r->m);
        if (ticket >= r->next_ticket) {
            g_runner_err = "pf_runner_wait: unknown ticket";
            return PF_ERR_INVALID_ARGUMENT;
        }
        r->cv_done.wait(lk, [&] { return r->launched >= ticket; });
        for (auto it = r->failures.begin(); it != r->failures.end(); ++it) {
            if (it->ticket == ticket) {
                const int rc = it->status;
                g_runner_err = it->what;
                r->failures.erase(it);
                return rc;
            }
        }
    }
    Slot &s = r->slots[(size_t)(ticket % (long long)r->slots.size())];
    cudaError_t e = cudaEventSynchronize(s.e1);
    if (e == cudaSuccess && device_ms && cudaEventElapsedTime(device_ms, s.e0, s.e1) != cudaSuccess) {
        cudaGetLastError();      // the slot is already being reused by a later batch: no timing for this one
        *device_ms = -1.f;
    }
    if (e != cudaSuccess) {
        g_runner_err = std::string("pf_runner_wait: ") + cudaGetErrorName(e);
        return (int)e;
    }
    return PF_OK;
}

int pf_runner_destroy(pf_runner *r)
{
    if (!r) return PF_OK;
    {
        std::lock_guard<std::mutex> lk(r->m);
        r->stop = true;
    }
    r->cv_job.notify_all();
    if (r->th.joinable()) r->th.join();
    for (auto &s : r->slots) {
        if (s.used) cudaEventSynchronize(s.e1);
        if (s.e0) cudaEventDestroy(s.e0);
        if (s.e1) cudaEventDestroy(s.e1);
    }
    delete r;
    return PF_OK;
}

}  // extern "C"
