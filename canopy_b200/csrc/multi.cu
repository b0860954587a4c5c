// Single-process multi-GPU run: contiguous Philox counter ranges per device, one host
// thread per device, then ONE ncclAllReduce(ncclUint64, ncclSum) over the final tallies on NVLink.
// The tallies buffer the engine finalises into is the buffer NCCL reduces in place on the same stream: no staging
// copy, and the per-device runs are enqueued without a report, so nothing synchronises the host between the last
// gate kernel and the collective (device times come from events read after the collective has been enqueued).
// Communicators are created once per device list and kept for the life of the process (never destroyed: NCCL
// teardown at exit is unsafe once the CUDA context is going away); a per-list mutex serialises the group calls.
//
// NCCL is bound lazily with dlopen("libnccl.so.2") so that libcanopy has no load-time dependency on it
// (the torchrun path lets torch.distributed own the communicator and only needs canopy_mc_run_device).
#include <dlfcn.h>
#include <nccl.h>  // types only; no symbol is linked

#include <chrono>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include <cuda_runtime.h>

#include "plan.h"

namespace canopy {
namespace {

struct NcclApi {
    void* h = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
};

NcclApi& nccl() {
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        api.h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!api.h) api.h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (!api.h) return;
        api.CommInitAll = (decltype(api.CommInitAll))dlsym(api.h, "ncclCommInitAll");
        api.CommDestroy = (decltype(api.CommDestroy))dlsym(api.h, "ncclCommDestroy");
        api.AllReduce = (decltype(api.AllReduce))dlsym(api.h, "ncclAllReduce");
        api.GroupStart = (decltype(api.GroupStart))dlsym(api.h, "ncclGroupStart");
        api.GroupEnd = (decltype(api.GroupEnd))dlsym(api.h, "ncclGroupEnd");
        api.GetErrorString = (decltype(api.GetErrorString))dlsym(api.h, "ncclGetErrorString");
    });
    if (!api.h || !api.CommInitAll || !api.AllReduce || !api.CommDestroy || !api.GroupStart || !api.GroupEnd)
        throw Error(CANOPY_E_NCCL, "libnccl.so.2 could not be loaded");
    return api;
}

void nccl_check(ncclResult_t r, const char* what) {
    if (r != ncclSuccess) {
        NcclApi& a = nccl();
        throw Error(CANOPY_E_NCCL, std::string(what) + ": " + (a.GetErrorString ? a.GetErrorString(r) : "nccl error"));
    }
}

struct CommSet { std::vector<ncclComm_t> comms; std::mutex mu; };
std::mutex g_comm_mu;
std::map<std::vector<int>, CommSet> g_comms;  // communicators are cached per device list for the process lifetime

CommSet& get_comms(const std::vector<int>& devs) {
    std::lock_guard<std::mutex> lock(g_comm_mu);
    auto it = g_comms.find(devs);
    if (it != g_comms.end()) return it->second;
    CommSet& cs = g_comms[devs];   // std::map: the reference stays valid for the life of the process
    cs.comms.resize(devs.size());
    try {
        nccl_check(nccl().CommInitAll(cs.comms.data(), (int)devs.size(), devs.data()), "ncclCommInitAll");
    } catch (...) { g_comms.erase(devs); throw; }
    return cs;
}

#define CU(expr) do { cudaError_t _e = (expr); if (_e != cudaSuccess) throw Error(CANOPY_E_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e)); } while (0)

}  // namespace

void mc_run_multi(canopy_plan* plan, const canopy_mc_opts& o, const int32_t* devices, int32_t n_devices,
                  uint64_t* tallies_host, canopy_mc_report* rep) {
    const auto t_begin = std::chrono::steady_clock::now();
    const size_t nt = plan->low.tally_kind.size();
    std::vector<int> devs(devices, devices + n_devices);
    if (n_devices == 1) {  // degenerate: no collective
        canopy_mc_opts o1 = o; o1.device = devs[0];
        mc_run(plan, o1, tallies_host, nullptr, rep);
        return;
    }
    CommSet& cs = get_comms(devs);
    std::vector<unsigned long long*> d_t(n_devices, nullptr);
    std::vector<cudaStream_t> st(n_devices, nullptr);
    std::vector<std::string> errs(n_devices);
    std::vector<canopy_status> codes(n_devices, CANOPY_OK);
    std::vector<std::thread> th;
    for (int i = 0; i < n_devices; ++i) {
        th.emplace_back([&, i] {
            try {
                CU(cudaSetDevice(devs[i]));
                CU(cudaMalloc(&d_t[i], std::max<size_t>(nt, 1) * 8));
                CU(cudaStreamCreateWithFlags(&st[i], cudaStreamNonBlocking));
                canopy_mc_opts oi = o;
                oi.device = devs[i]; oi.stream = st[i];
                shard_range(o.sample_begin, o.n_samples, i, n_devices, &oi.sample_begin, &oi.n_samples);
                mc_run(plan, oi, nullptr, (uint64_t*)d_t[i], nullptr);   // no report: returns with the kernels in flight
            } catch (const Error& e) { codes[i] = e.code; errs[i] = e.what(); }
            catch (const std::exception& e) { codes[i] = CANOPY_E_OOM; errs[i] = e.what(); }   // bad_alloc, system_error: never std::terminate
            catch (...) { codes[i] = CANOPY_E_INVALID; errs[i] = "unknown error in a device worker"; }
        });
    }
    for (auto& t : th) t.join();
    auto cleanup = [&] {
        for (int i = 0; i < n_devices; ++i) {
            cudaSetDevice(devs[i]);
            if (st[i]) { cudaStreamSynchronize(st[i]); cudaStreamDestroy(st[i]); }
            cudaFree(d_t[i]);
        }
    };
    for (int i = 0; i < n_devices; ++i)
        if (codes[i] != CANOPY_OK) { cleanup(); throw Error(codes[i], "device " + std::to_string(devs[i]) + ": " + errs[i]); }
    try {
        if (nt) {
            std::lock_guard<std::mutex> comm_lock(cs.mu);   // two concurrent runs on one device list must not interleave their groups
            nccl_check(nccl().GroupStart(), "ncclGroupStart");
            try {
                for (int i = 0; i < n_devices; ++i)
                    nccl_check(nccl().AllReduce(d_t[i], d_t[i], nt, ncclUint64, ncclSum, cs.comms[i], st[i]), "ncclAllReduce");
            } catch (...) { nccl().GroupEnd(); throw; }   // never leave the group open: the next NCCL call of this thread would join it
            nccl_check(nccl().GroupEnd(), "ncclGroupEnd");
            CU(cudaSetDevice(devs[0]));
            CU(cudaMemcpyAsync(tallies_host, d_t[0], nt * 8, cudaMemcpyDeviceToHost, st[0]));
        }
        for (int i = 0; i < n_devices; ++i) { CU(cudaSetDevice(devs[i])); CU(cudaStreamSynchronize(st[i])); }
    } catch (...) { cleanup(); throw; }
    cleanup();
    if (rep) {   // wall-clock of the whole call on the host; per-kernel times need a single-device run with profile = 1
        *rep = canopy_mc_report{};
        rep->total_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_begin).count();
        rep->d2h_bytes = nt * 8;
    }
}

}  // namespace canopy
