// C ABI, part 4: the two exchange steps of the multi-GPU paths as entry points a C++ server can call with its own
// ncclComm_t -- no Python, no torch:
//   icx_scores_allgather        config 4: ONE all-gather of {flagged, incongruent, score} per world-line candidate
//   icx_timeline_slot_send/recv config 5: the future -> past snapshot stream between timeslices on different GPUs
// NCCL is bound at run time (dlopen / dlsym of libnccl.so.2): the library has no link-time dependency on it, and inside a
// process that already carries an NCCL (torch's bundled one) exactly that copy is used.  Without NCCL: ICX_E_STATE.
#include <dlfcn.h>
#include <cstring>
#include <string>
#include <vector>
#include "icx_host.h"
#include "../../include/icx_nccl.h"

using namespace icx;

// accessors implemented in icx_timeline.cu (IcxTimeline is opaque here)
int icx_tl_scan_buffers(IcxTimeline* tl, uint32_t n_worldlines, const uint32_t** d_count, const unsigned long long** d_score, const uint8_t** d_ever,
                        cudaStream_t* scan_stream, int* device);
void icx_tl_set_error(IcxTimeline* tl, const char* msg);
int icx_tl_device(const IcxTimeline* tl);

namespace {

typedef int (*fn_allgather)(const void*, void*, size_t, int, void*, cudaStream_t);
typedef int (*fn_sendrecv)(void*, size_t, int, int, void*, cudaStream_t);
typedef int (*fn_void)(void);
typedef const char* (*fn_errstr)(int);
struct Nccl { void* h = nullptr; fn_allgather all_gather = nullptr; fn_sendrecv send = nullptr, recv = nullptr; fn_void group_start = nullptr, group_end = nullptr;
              fn_errstr err = nullptr; bool tried = false; };
Nccl g_nccl;

bool nccl_bind()
{
    if (g_nccl.tried) return g_nccl.all_gather != nullptr;
    g_nccl.tried = true;
    void* h = dlopen(nullptr, RTLD_NOW);                                   // an NCCL already in the process (e.g. torch's)
    if (!h || !dlsym(h, "ncclAllGather")) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) return false;
    g_nccl.h = h;
    g_nccl.all_gather = (fn_allgather)dlsym(h, "ncclAllGather");
    g_nccl.send = (fn_sendrecv)dlsym(h, "ncclSend"); g_nccl.recv = (fn_sendrecv)dlsym(h, "ncclRecv");
    g_nccl.group_start = (fn_void)dlsym(h, "ncclGroupStart"); g_nccl.group_end = (fn_void)dlsym(h, "ncclGroupEnd");
    g_nccl.err = (fn_errstr)dlsym(h, "ncclGetErrorString");
    return g_nccl.all_gather && g_nccl.send && g_nccl.recv && g_nccl.group_start && g_nccl.group_end;
}
const int NCCL_UINT8 = 1;                                                   // ncclUint8 / ncclChar family: ncclInt8 = 0, ncclUint8 = 1

struct Packed { uint32_t ever, count; unsigned long long score; };          // 16 B per world-line
__global__ void k_pack_scores(const uint32_t* __restrict__ count, const unsigned long long* __restrict__ score, const uint8_t* __restrict__ ever,
                              uint32_t n, uint32_t n_pad, Packed* __restrict__ out)
{
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n_pad) return;
    Packed p; p.ever = 0; p.count = 0; p.score = 0;
    if (k < n) { p.ever = ever[k]; p.count = count[k]; p.score = score[k]; }
    out[k] = p;
}

DeviceBuffer g_send[64], g_recv[64];                                       // per device staging of the gather

// NCCL launches the operations of a group at ncclGroupEnd: an import enqueued right after ncclRecv inside a group would run
// BEFORE the receive.  Receives posted inside icx_nccl_group_start / _end are therefore imported by icx_nccl_group_end.
struct PendingImport { IcxTimeline* tl; uint16_t coherency; uint32_t n; void* scratch; size_t bytes; void* stream; };
thread_local int g_group_depth = 0;
thread_local std::vector<PendingImport> g_pending;

} // namespace

extern "C" {

int icx_nccl_available(void) { return nccl_bind() ? 1 : 0; }

int icx_scores_allgather(IcxTimeline* tl, uint32_t n_local, uint32_t n_per_rank, int n_ranks, ncclComm_t comm,
                         uint8_t* flags_ever, uint32_t* incongruent, double* score)
{
    if (!tl || !comm || n_ranks <= 0 || n_local > n_per_rank) return ICX_E_INVALID;
    if (!nccl_bind()) { icx_tl_set_error(tl, "scores_allgather: NCCL (libnccl.so.2) is not available in this process"); return ICX_E_STATE; }
    const uint32_t* d_count; const unsigned long long* d_score; const uint8_t* d_ever; cudaStream_t st; int device;
    int rc = icx_tl_scan_buffers(tl, n_local, &d_count, &d_score, &d_ever, &st, &device);
    if (rc != ICX_OK) return rc;
    cudaSetDevice(device);
    DeviceBuffer& sb = g_send[device & 63]; DeviceBuffer& rb = g_recv[device & 63];
    if (sb.ensure((size_t)n_per_rank * sizeof(Packed)) != cudaSuccess || rb.ensure((size_t)n_per_rank * n_ranks * sizeof(Packed)) != cudaSuccess) return ICX_E_NOMEM;
    k_pack_scores<<<(n_per_rank + 255) / 256, 256, 0, st>>>(d_count, d_score, d_ever, n_local, n_per_rank, (Packed*)sb.ptr);
    int nrc = g_nccl.all_gather(sb.ptr, rb.ptr, (size_t)n_per_rank * sizeof(Packed), NCCL_UINT8, comm, st);
    if (nrc != 0) { icx_tl_set_error(tl, (std::string("ncclAllGather: ") + (g_nccl.err ? g_nccl.err(nrc) : "error")).c_str()); return ICX_E_CUDA; }
    std::vector<Packed> h((size_t)n_per_rank * n_ranks);
    if (cudaMemcpyAsync(h.data(), rb.ptr, h.size() * sizeof(Packed), cudaMemcpyDeviceToHost, st) != cudaSuccess || cudaStreamSynchronize(st) != cudaSuccess) {
        icx_tl_set_error(tl, "scores_allgather: copy back failed"); return ICX_E_CUDA;
    }
    for (size_t k = 0; k < h.size(); k++) {                                 // rank-major: [rank][slot]
        if (flags_ever) flags_ever[k] = h[k].ever ? 1 : 0;
        if (incongruent) incongruent[k] = h[k].count;
        if (score) score[k] = (double)h[k].score / 65536.0;
    }
    return ICX_OK;
}

int icx_timeline_slot_send(IcxTimeline* tl, uint16_t coherency, ncclComm_t comm, int peer, void* device_scratch, size_t scratch_bytes, void* stream)
{
    if (!tl || !comm || !device_scratch) return ICX_E_INVALID;
    if (!nccl_bind()) { icx_tl_set_error(tl, "slot_send: NCCL is not available in this process"); return ICX_E_STATE; }
    uint32_t n = 0;
    int rc = icx_timeline_slot_export(tl, coherency, device_scratch, scratch_bytes, &n, stream);
    if (rc != ICX_OK) return rc;
    cudaSetDevice(icx_tl_device(tl));
    int nrc = g_nccl.send(device_scratch, scratch_bytes, NCCL_UINT8, peer, comm, (cudaStream_t)stream);      // fixed size: both sides agree on the capacity
    if (nrc != 0) { icx_tl_set_error(tl, (std::string("ncclSend: ") + (g_nccl.err ? g_nccl.err(nrc) : "error")).c_str()); return ICX_E_CUDA; }
    return ICX_OK;
}

int icx_timeline_slot_recv(IcxTimeline* tl, uint16_t coherency, uint32_t n_entities, ncclComm_t comm, int peer, void* device_scratch, size_t scratch_bytes, void* stream)
{
    if (!tl || !comm || !device_scratch) return ICX_E_INVALID;
    if (!nccl_bind()) { icx_tl_set_error(tl, "slot_recv: NCCL is not available in this process"); return ICX_E_STATE; }
    cudaSetDevice(icx_tl_device(tl));
    int nrc = g_nccl.recv(device_scratch, scratch_bytes, NCCL_UINT8, peer, comm, (cudaStream_t)stream);
    if (nrc != 0) { icx_tl_set_error(tl, (std::string("ncclRecv: ") + (g_nccl.err ? g_nccl.err(nrc) : "error")).c_str()); return ICX_E_CUDA; }
    if (g_group_depth > 0) { g_pending.push_back({tl, coherency, n_entities, device_scratch, scratch_bytes, stream}); return ICX_OK; }
    return icx_timeline_slot_import(tl, coherency, n_entities, device_scratch, scratch_bytes, stream);
}

int icx_nccl_group_start(void)
{
    if (!nccl_bind() || g_nccl.group_start() != 0) return ICX_E_STATE;
    g_group_depth++;
    return ICX_OK;
}
int icx_nccl_group_end(void)
{
    if (!nccl_bind() || g_group_depth <= 0) return ICX_E_STATE;
    const int nrc = g_nccl.group_end();
    g_group_depth--;
    if (nrc != 0) { g_pending.clear(); return ICX_E_CUDA; }
    int rc = ICX_OK;
    if (g_group_depth == 0) {
        for (const PendingImport& p : g_pending) { const int r = icx_timeline_slot_import(p.tl, p.coherency, p.n, p.scratch, p.bytes, p.stream); if (r != ICX_OK) rc = r; }
        g_pending.clear();
    }
    return rc;
}

} // extern "C"
