// dist.cpp — the one exchange step of the sharded iteration: a sum-allreduce of the
// packed buffer [numerator | row sums of H] over NCCL (NVLink 5 / NVSwitch), one process per GPU.
// NCCL is dlopen'ed so that the single-GPU path has no dependency on it; inside a torch process
// the soname resolves to the libnccl torch already loaded.
#include <dlfcn.h>
#include <atomic>
#include <cstring>
#include <vector>

#include "session.hpp"

namespace nmf {

namespace {
typedef void* ncclComm_t;
struct ncclUniqueId { char internal[128]; };
enum { ncclSum = 0, ncclFloat32 = 7, ncclFloat64 = 8 };

struct Api {
    void* h = nullptr;
    int (*GetUniqueId)(ncclUniqueId*) = nullptr;
    int (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*CommDestroy)(ncclComm_t) = nullptr;
    int (*CommAbort)(ncclComm_t) = nullptr;
    int (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
};

Api& api() {
    static Api a;
    if (a.h) return a;
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* n : names) { a.h = dlopen(n, RTLD_NOW | RTLD_GLOBAL); if (a.h) break; }
    NMF_CHECK(a.h != nullptr, std::string("cannot load NCCL (libnccl.so.2): ") + (dlerror() ? dlerror() : ""));
    auto sym = [&](const char* s) { void* p = dlsym(a.h, s); NMF_CHECK(p != nullptr, std::string("NCCL symbol missing: ") + s); return p; };
    a.GetUniqueId    = (int (*)(ncclUniqueId*))sym("ncclGetUniqueId");
    a.CommInitRank   = (int (*)(ncclComm_t*, int, ncclUniqueId, int))sym("ncclCommInitRank");
    a.AllReduce      = (int (*)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t))sym("ncclAllReduce");
    a.CommDestroy    = (int (*)(ncclComm_t))sym("ncclCommDestroy");
    a.CommAbort      = (int (*)(ncclComm_t))sym("ncclCommAbort");
    a.CommInitAll    = (int (*)(ncclComm_t*, int, const int*))sym("ncclCommInitAll");
    a.GetErrorString = (const char* (*)(int))sym("ncclGetErrorString");
    return a;
}

void check(int r, const char* what) {
    if (r != 0) NMF_FAIL(std::string(what) + ": " + api().GetErrorString(r));
}
}  // namespace

struct Nccl { ncclComm_t comm = nullptr; int refs = 1; std::atomic<bool> aborted{false}; };   // shared by the sessions of one process (nccl_share)

int nccl_unique_id(void* id128) {
    ncclUniqueId id;
    check(api().GetUniqueId(&id), "ncclGetUniqueId");
    std::memcpy(id128, &id, sizeof(id));
    return 0;
}

void nccl_join(nmf_session_impl& s, int rank, int nranks, const void* id128) {
    NMF_CHECK(nranks >= 1 && rank >= 0 && rank < nranks, "bad rank / nranks");
    if (nranks == 1) { s.rank = 0; s.nranks = 1; return; }
    NMF_CHECK(s.g_full == nullptr, "join the communicator before the first run");
    CUDA_CHECK(cudaSetDevice(s.device));
    ncclUniqueId id;
    std::memcpy(&id, id128, sizeof(id));
    auto* n = new Nccl();
    const int r = api().CommInitRank(&n->comm, nranks, id, rank);
    if (r != 0) { delete n; check(r, "ncclCommInitRank"); }
    s.nccl = n; s.rank = rank; s.nranks = nranks;
}

// One process, a host thread per GPU (update_div with NMF_NUM_GPUS = g): all g communicators in ONE call from the
// calling thread — no rank can be left waiting in ncclCommInitRank for a peer that failed before it got there.
void nccl_init_all(nmf_session_impl* const* sessions, int g) {
    if (g <= 1) { if (g == 1) { sessions[0]->rank = 0; sessions[0]->nranks = 1; } return; }
    std::vector<int> devs((size_t)g);
    std::vector<ncclComm_t> comms((size_t)g, nullptr);
    for (int r = 0; r < g; ++r) {
        NMF_CHECK(sessions[r]->g_full == nullptr && sessions[r]->nccl == nullptr, "join the communicator before the first run, once");
        devs[(size_t)r] = sessions[r]->device;
    }
    check(api().CommInitAll(comms.data(), g, devs.data()), "ncclCommInitAll");
    for (int r = 0; r < g; ++r) {
        auto* n = new Nccl();
        n->comm = comms[(size_t)r];
        sessions[r]->nccl = n; sessions[r]->rank = r; sessions[r]->nranks = g;
    }
}

// A rank that failed inside the iteration calls this on EVERY session of the process: peers blocked in a collective
// that will never complete return with an error instead of hanging the call.
void nccl_abort(nmf_session_impl& s) {
    Nccl* n = s.nccl;
    if (!n || !n->comm) return;
    if (n->aborted.exchange(true)) return;
    api().CommAbort(n->comm);
}

void nccl_allreduce_f32(nmf_session_impl& s, float* buf, size_t n) {
    NMF_CHECK(s.nccl != nullptr, "allreduce without a communicator");
    check(api().AllReduce(buf, buf, n, ncclFloat32, ncclSum, s.nccl->comm, s.stream), "ncclAllReduce(f32)");
}

void nccl_allreduce_f64(nmf_session_impl& s, double* buf, size_t n) {
    NMF_CHECK(s.nccl != nullptr, "allreduce without a communicator");
    check(api().AllReduce(buf, buf, n, ncclFloat64, ncclSum, s.nccl->comm, s.stream), "ncclAllReduce(f64)");
}

// A second session of the same process (same rank, same device) reuses the communicator of `src`:
// ncclCommInitRank costs seconds, a process makes it once.
void nccl_share(nmf_session_impl& dst, nmf_session_impl& src) {
    NMF_CHECK(dst.nccl == nullptr && dst.g_full == nullptr, "share the communicator before the first run, once");
    NMF_CHECK(dst.device == src.device, "sessions that share a communicator must be on the same device");
    if (src.nccl == nullptr) { dst.rank = 0; dst.nranks = 1; return; }
    src.nccl->refs += 1;
    dst.nccl = src.nccl; dst.rank = src.rank; dst.nranks = src.nranks;
}

void nccl_destroy(Nccl* n) {
    if (!n) return;
    if (--n->refs > 0) return;
    if (n->comm && !n->aborted.load()) api().CommDestroy(n->comm);     // ncclCommAbort already freed an aborted communicator
    delete n;
}

}  // namespace nmf
