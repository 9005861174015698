// sfp_multi.cu — batched independent worlds over SEVERAL GPUs from ONE process (sfp.h,
// sfp_batch_multi_*): the engine is a single process, so the library itself spreads a batch
// over the devices it is given.  One SfpBatch and one worker thread per device; the world list
// is cut into contiguous blocks (device d owns worlds [first[d], first[d + 1])); every call fans
// out to the workers and joins them.  There is no exchange between worlds, hence no collective:
// per-world statistics are simply concatenated on the host (north_star: "NCCL used only to
// gather statistics" — inside one process a gather is a memcpy).
//
// Reference: none (the reference is a single-GPU-less CPU engine; BASELINE.json config 5).
#include <cuda_runtime.h>

#include <algorithm>
#include <condition_variable>
#include <cstring>
#include <functional>
#include <mutex>
#include <new>
#include <string>
#include <thread>
#include <vector>

#include "sfp.h"

namespace {

// one worker per device: runs the jobs it is handed, in order
class Worker {
   public:
    Worker() : th_([this] { loop(); }) {}
    ~Worker() {
        {
            std::lock_guard<std::mutex> lk(m_);
            stop_ = true;
        }
        cv_.notify_all();
        th_.join();
    }
    void submit(std::function<int32_t()> job) {
        {
            std::lock_guard<std::mutex> lk(m_);
            job_ = std::move(job);
            busy_ = true;
        }
        cv_.notify_all();
    }
    int32_t wait() {
        std::unique_lock<std::mutex> lk(m_);
        done_cv_.wait(lk, [&] { return !busy_; });
        return rc_;
    }

   private:
    void loop() {
        for (;;) {
            std::function<int32_t()> job;
            {
                std::unique_lock<std::mutex> lk(m_);
                cv_.wait(lk, [&] { return stop_ || (busy_ && job_); });
                if (stop_) return;
                job = std::move(job_);
                job_ = nullptr;
            }
            int32_t rc = job();
            {
                std::lock_guard<std::mutex> lk(m_);
                rc_ = rc;
                busy_ = false;
            }
            done_cv_.notify_all();
        }
    }
    std::mutex m_;
    std::condition_variable cv_, done_cv_;
    std::function<int32_t()> job_;
    bool busy_ = false, stop_ = false;
    int32_t rc_ = SFP_OK;
    std::thread th_;  // last: the thread starts when everything above exists
};

}  // namespace

struct SfpBatchMulti {
    uint32_t n_worlds = 0, bpw = 0;
    std::vector<int32_t> devices;
    std::vector<uint32_t> first;  // n_dev + 1 entries
    std::vector<SfpBatch*> parts;
    std::vector<Worker*> workers;
    std::string err;

    // job(d) on every device's worker, concurrently; the first error wins
    int32_t fan_out(const std::function<int32_t(uint32_t)>& job) {
        for (uint32_t d = 0; d < parts.size(); ++d) workers[d]->submit([&job, d] { return job(d); });
        int32_t rc = SFP_OK;
        for (uint32_t d = 0; d < parts.size(); ++d) {
            int32_t r = workers[d]->wait();
            if (r != SFP_OK && rc == SFP_OK) {
                rc = r;
                err = "device " + std::to_string(devices[d]) + ": " + sfp_batch_last_error(parts[d]);
            }
        }
        return rc;
    }
    size_t body0(uint32_t d) const { return (size_t)first[d] * bpw; }
    uint32_t worlds(uint32_t d) const { return first[d + 1] - first[d]; }
};

extern "C" {

int32_t sfp_batch_create_multi(const SfpParams* params, uint32_t n_worlds, uint32_t bodies_per_world,
                               const int32_t* device_ids, uint32_t n_dev, SfpBatchMulti** out) {
    if (!out) return SFP_E_INVALID;
    *out = nullptr;
    if (!device_ids || n_dev == 0 || n_dev > 64 || n_worlds < n_dev || bodies_per_world == 0)
        return SFP_E_INVALID;
    SfpParams p;
    if (params) p = *params;
    else sfp_default_params(&p);
    SfpBatchMulti* m = new (std::nothrow) SfpBatchMulti();
    if (!m) return SFP_E_NOMEM;
    m->n_worlds = n_worlds;
    m->bpw = bodies_per_world;
    m->devices.assign(device_ids, device_ids + n_dev);
    m->first.resize(n_dev + 1);
    const uint32_t base = n_worlds / n_dev, extra = n_worlds % n_dev;
    for (uint32_t d = 0; d <= n_dev; ++d) m->first[d] = d * base + std::min(d, extra);
    m->parts.assign(n_dev, nullptr);
    for (uint32_t d = 0; d < n_dev; ++d) m->workers.push_back(new Worker());
    int32_t rc = m->fan_out([&](uint32_t d) {
        SfpParams pd = p;
        pd.device = m->devices[d];
        return sfp_batch_create(&pd, m->worlds(d), bodies_per_world, &m->parts[d]);
    });
    if (rc != SFP_OK) {
        sfp_batch_multi_destroy(m);
        return rc;
    }
    *out = m;
    return SFP_OK;
}

int32_t sfp_batch_multi_destroy(SfpBatchMulti* m) {
    if (!m) return SFP_OK;
    if (!m->workers.empty())
        m->fan_out([&](uint32_t d) { return m->parts[d] ? sfp_batch_destroy(m->parts[d]) : (int32_t)SFP_OK; });
    for (Worker* w : m->workers) delete w;
    delete m;
    return SFP_OK;
}

int32_t sfp_batch_multi_partition(SfpBatchMulti* m, uint32_t dev_index, int32_t* device,
                                  uint32_t* first_world, uint32_t* n_worlds) {
    if (!m || dev_index >= m->parts.size()) return SFP_E_INVALID;
    if (device) *device = m->devices[dev_index];
    if (first_world) *first_world = m->first[dev_index];
    if (n_worlds) *n_worlds = m->worlds(dev_index);
    return SFP_OK;
}

int32_t sfp_batch_multi_upload(SfpBatchMulti* m, const SfpBodyDesc* bodies) {
    if (!m || !bodies) return SFP_E_INVALID;
    return m->fan_out([&](uint32_t d) { return sfp_batch_upload(m->parts[d], bodies + m->body0(d)); });
}

int32_t sfp_batch_multi_set_polygons(SfpBatchMulti* m, const SfpPolygon* polys, uint32_t n) {
    if (!m) return SFP_E_INVALID;
    return m->fan_out([&](uint32_t d) { return sfp_batch_set_polygons(m->parts[d], polys, n); });
}

int32_t sfp_batch_multi_set_joints(SfpBatchMulti* m, const SfpJointDesc* joints,
                                   uint32_t joints_per_world, const uint32_t* counts) {
    if (!m) return SFP_E_INVALID;
    return m->fan_out([&](uint32_t d) {
        return sfp_batch_set_joints(m->parts[d],
                                    joints ? joints + (size_t)m->first[d] * joints_per_world : nullptr,
                                    joints_per_world, counts ? counts + m->first[d] : nullptr);
    });
}

int32_t sfp_batch_multi_step(SfpBatchMulti* m, float dt, uint32_t substeps) {
    if (!m) return SFP_E_INVALID;
    return m->fan_out([&](uint32_t d) { return sfp_batch_step(m->parts[d], dt, substeps); });
}

int32_t sfp_batch_multi_sync(SfpBatchMulti* m) {
    if (!m) return SFP_E_INVALID;
    return m->fan_out([&](uint32_t d) { return sfp_batch_sync(m->parts[d]); });
}

int32_t sfp_batch_multi_download(SfpBatchMulti* m, SfpBodyState* out) {
    if (!m || !out) return SFP_E_INVALID;
    return m->fan_out([&](uint32_t d) { return sfp_batch_download(m->parts[d], out + m->body0(d)); });
}

int32_t sfp_batch_multi_step_host(SfpBatchMulti* m, const SfpBodyState* in, SfpBodyState* out, float dt,
                                  uint32_t substeps) {
    if (!m || !in || !out) return SFP_E_INVALID;
    return m->fan_out([&](uint32_t d) {
        return sfp_batch_step_host(m->parts[d], in + m->body0(d), out + m->body0(d), dt, substeps);
    });
}

int32_t sfp_batch_multi_stats(SfpBatchMulti* m, SfpStats* out) {
    if (!m || !out) return SFP_E_INVALID;
    return m->fan_out([&](uint32_t d) { return sfp_batch_stats(m->parts[d], out + m->first[d]); });
}

int32_t sfp_batch_multi_last_step_ms(SfpBatchMulti* m, float* ms) {
    if (!m || !ms) return SFP_E_INVALID;
    std::vector<float> t(m->parts.size(), 0.0f);
    int32_t rc = m->fan_out([&](uint32_t d) { return sfp_batch_last_step_ms(m->parts[d], &t[d]); });
    *ms = 0.0f;
    for (float v : t) *ms = std::max(*ms, v);  // the slowest device is the step
    return rc;
}

const char* sfp_batch_multi_last_error(SfpBatchMulti* m) { return m ? m->err.c_str() : "null batch"; }

}  // extern "C"
