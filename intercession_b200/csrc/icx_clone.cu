// C ABI, part 2: ParallelCosmosContext::load_and_link / extract_entity_updates
// (source/spacetime/parallel_cosmos_context.cpp:60-169, :171-423) between two device worlds.
// The reference deep-copies a cosmos entity by entity through serialisation; here the rows move
// device to device (same GPU or peer GPU over NVLink), the host only does the id bookkeeping.
#include <algorithm>
#include <cstring>
#include <unordered_map>
#include "icx_host.h"

using namespace icx;

namespace {

int cfail(IcxWorld* w, int code, const std::string& msg) { if (w) w->error = msg; return code; }
int cfail_cuda(IcxWorld* w, cudaError_t e, const char* what)
{
    if (w) w->error = std::string(what) + ": " + cudaGetErrorString(e);
    return ICX_E_CUDA;
}
#define CKW(w, call) do { cudaError_t _e = (call); if (_e != cudaSuccess) return cfail_cuda(w, _e, #call); } while (0)

Mirror mirror(IcxWorld* w)
{
    Mirror m;
    m.ids = (uint32_t*)w->mir_ids.ptr; m.T = (IcxTransform*)w->mir_T.ptr; m.P = (IcxPhysics*)w->mir_P.ptr;
    m.C = (IcxColliderComponent*)w->mir_C.ptr; m.hasP = (uint8_t*)w->mir_hasP.ptr;
    return m;
}

// src rows idx[0..count) -> compact staging on the source device (AoS mirror refreshed first)
int gather(IcxWorld* src, const std::vector<uint32_t>& idx, int mode)
{
    const uint32_t count = (uint32_t)idx.size(), n = src->n_entities;
    cudaSetDevice(src->device);
    cudaStream_t st = src->stream;
    CKW(src, src->cl_idx.ensure((size_t)std::max(count, 1u) * 4)); CKW(src, src->cl_ids.ensure((size_t)std::max(count, 1u) * 4));
    CKW(src, src->cl_T.ensure((size_t)std::max(count, 1u) * sizeof(IcxTransform))); CKW(src, src->cl_hasP.ensure(std::max(count, 1u)));
    CKW(src, src->cl_ts.ensure(std::max(count, 1u))); CKW(src, src->cl_tc.ensure((size_t)std::max(count, 1u) * 2));
    if (src->have_P) CKW(src, src->cl_P.ensure((size_t)std::max(count, 1u) * sizeof(IcxPhysics)));
    if (src->have_C) CKW(src, src->cl_C.ensure((size_t)std::max(count, 1u) * sizeof(IcxColliderComponent)));
    launch_export_kernel(src->W, mirror(src), 1, src->have_P ? 1 : 0, src->have_C ? 1 : 0, nullptr, nullptr, n, st);
    if (count) CKW(src, cudaMemcpyAsync(src->cl_idx.ptr, idx.data(), (size_t)count * 4, cudaMemcpyHostToDevice, st));
    launch_gather_rows(src->W, mirror(src), (const uint32_t*)src->cl_idx.ptr, count, mode, (uint32_t*)src->cl_ids.ptr,
                       (IcxTransform*)src->cl_T.ptr, src->have_P ? (IcxPhysics*)src->cl_P.ptr : nullptr,
                       src->have_C ? (IcxColliderComponent*)src->cl_C.ptr : nullptr, (uint8_t*)src->cl_hasP.ptr,
                       (uint8_t*)src->cl_ts.ptr, (uint16_t*)src->cl_tc.ptr, st);
    CKW(src, cudaStreamSynchronize(st));
    CKW(src, cudaGetLastError());
    return ICX_OK;
}

// device-to-device copy between the two worlds' devices on the destination's stream
cudaError_t xcopy(IcxWorld* dst, void* d, IcxWorld* src, const void* s, size_t bytes)
{
    if (!bytes) return cudaSuccess;
    if (dst->device == src->device) return cudaMemcpyAsync(d, s, bytes, cudaMemcpyDeviceToDevice, dst->stream);
    return cudaMemcpyPeerAsync(d, dst->device, s, src->device, bytes, dst->stream);
}

} // namespace

extern "C" {

int icx_world_clone(IcxWorld* src, IcxWorld* dst, const uint8_t* has_future, uint32_t* n_cloned)
{
    if (!src || !dst || src == dst) return ICX_E_INVALID;
    const uint32_t n = src->n_entities;
    // :110-116 omit what does not exist in the future at all: chain-link 0 without a future timestream
    std::vector<uint32_t> idx; idx.reserve(n);
    for (uint32_t i = 0; i < n; i++)
        if (!(has_future && !has_future[i] && (src->h_ids[i] & 0xFu) == 0u)) idx.push_back(i);
    const uint32_t nk = (uint32_t)idx.size();
    { int mrc = icx_host_masks(src); if (mrc != ICX_OK) { dst->error = src->error; return mrc; } }
    if (nk > dst->cap_entities) return cfail(dst, ICX_E_CAPACITY, "clone: destination max_entities too small");
    std::vector<uint32_t> first(nk);
    dst->active_mask.resize(nk); dst->h_ids.resize(nk);
    uint32_t np = 0;
    for (uint32_t i = 0; i < nk; i++) {
        const uint8_t m = src->active_mask[idx[i]];
        const uint32_t cnt = (m & 1u) + ((m >> 1) & 1u);
        dst->active_mask[i] = m; dst->h_ids[i] = src->h_ids[idx[i]];
        first[i] = np | (cnt << 30); np += cnt;
    }
    if (np > dst->cap_packets) return cfail(dst, ICX_E_CAPACITY, "clone: too many active colliders for the destination");
    int rc = gather(src, idx, 0);
    if (rc != ICX_OK) { dst->error = src->error; return rc; }

    cudaSetDevice(dst->device);
    cudaStream_t st = dst->stream;
    DeviceWorld& W = dst->W;
    const size_t cap = dst->cap_entities;
    CKW(dst, dst->mir_ids.ensure(cap * 4)); CKW(dst, dst->mir_T.ensure(cap * sizeof(IcxTransform))); CKW(dst, dst->mir_hasP.ensure(cap));
    if (src->have_P) CKW(dst, dst->mir_P.ensure(cap * sizeof(IcxPhysics)));
    if (src->have_C) CKW(dst, dst->mir_C.ensure(cap * sizeof(IcxColliderComponent)));
    CKW(dst, dst->st_ts.ensure(cap)); CKW(dst, dst->st_tc.ensure(cap * 2));
    dst->have_P = src->have_P; dst->have_C = src->have_C;
    CKW(dst, xcopy(dst, dst->mir_ids.ptr, src, src->cl_ids.ptr, (size_t)nk * 4));
    CKW(dst, xcopy(dst, dst->mir_T.ptr, src, src->cl_T.ptr, (size_t)nk * sizeof(IcxTransform)));
    CKW(dst, xcopy(dst, dst->mir_hasP.ptr, src, src->cl_hasP.ptr, nk));
    if (src->have_P) CKW(dst, xcopy(dst, dst->mir_P.ptr, src, src->cl_P.ptr, (size_t)nk * sizeof(IcxPhysics)));
    if (src->have_C) CKW(dst, xcopy(dst, dst->mir_C.ptr, src, src->cl_C.ptr, (size_t)nk * sizeof(IcxColliderComponent)));
    CKW(dst, xcopy(dst, dst->st_ts.ptr, src, src->cl_ts.ptr, nk));
    CKW(dst, xcopy(dst, dst->st_tc.ptr, src, src->cl_tc.ptr, (size_t)nk * 2));
    if (nk) CKW(dst, cudaMemcpyAsync(W.first_packet, first.data(), (size_t)nk * 4, cudaMemcpyHostToDevice, st));
    CKW(dst, dst->d_mask.ensure(cap));
    if (nk) CKW(dst, cudaMemcpyAsync(dst->d_mask.ptr, dst->active_mask.data(), nk, cudaMemcpyHostToDevice, st));
    dst->masks_on_host = true;
    CKW(dst, cudaMemsetAsync(W.counters, 0, CTR_TOTAL * sizeof(uint32_t), st));
    CKW(dst, cudaMemsetAsync(W.bm_keep, 0, (size_t)dst->cap_entities * sizeof(float), st)); CKW(dst, cudaMemsetAsync(W.bm_need, 0, (size_t)dst->cap_entities * sizeof(uint32_t), st));
    W.n_entities = nk; W.n_packets = np; dst->n_entities = nk; dst->n_packets = np;
    icx_size_buckets(dst);
    W.coherency = src->W.coherency;                                                 // :91 set_coherency(source)
    Mirror M = mirror(dst);
    if (!dst->have_P) M.P = nullptr;
    launch_import_kernel(W, M, nullptr, M.ids, M.T, M.P, dst->have_C ? M.C : nullptr, M.hasP, (const uint8_t*)dst->st_ts.ptr,
                         (const uint16_t*)dst->st_tc.ptr, nk, 1, dst->have_C ? 1 : 0, ++dst->c_epoch, st);
    dst->pend_any = false;
    CKW(dst, cudaMemcpyAsync(dst->h_counters, W.counters, CTR_TOTAL * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    CKW(dst, cudaStreamSynchronize(st));
    CKW(dst, cudaGetLastError());
    if (!(dst->cfg.cell_size > 0)) {                                                // same rule as icx_bodies_upload
        float ext; std::memcpy(&ext, &dst->h_counters[CTR_EXTENT], 4);
        if (!(ext > 0)) ext = 1.0f;
        W.cell_size = (ext + 2.0f * W.margin) * 1.02f; W.inv_cell = 1.0f / W.cell_size;
    }
    dst->stats.n_entities = nk; dst->stats.n_packets = np;
    dst->have_results = false;
    if (n_cloned) *n_cloned = nk;
    return ICX_OK;
}

int icx_world_extract(IcxWorld* par, IcxWorld* dst, uint32_t* n_extracted, uint32_t* n_missing)
{
    if (!par || !dst || par == dst) return ICX_E_INVALID;
    if (n_extracted) *n_extracted = 0;
    if (n_missing) *n_missing = 0;
    if (par->have_C && !dst->have_C) return cfail(dst, ICX_E_STATE, "extract: destination has no collider components");
    if (par->have_P && !dst->have_P) return cfail(dst, ICX_E_STATE, "extract: destination has no physics components");
    { int mrc = icx_host_masks(par); if (mrc == ICX_OK) mrc = icx_host_masks(dst); if (mrc != ICX_OK) return cfail(dst, mrc, "extract: cannot read the collider activation masks"); }
    // Both id lists are normally ascending (upload order = the reference's signature-map order, and a clone keeps it):
    // then id - 1 is matched by one merge pass; otherwise through a hash map.
    const bool merge = std::is_sorted(dst->h_ids.begin(), dst->h_ids.begin() + dst->n_entities) &&
                       std::is_sorted(par->h_ids.begin(), par->h_ids.begin() + par->n_entities);
    std::unordered_map<uint32_t, uint32_t> where;
    if (!merge) {
        where.reserve(dst->h_ids.size() * 2);
        for (uint32_t j = 0; j < dst->n_entities; j++) where.emplace(dst->h_ids[j], j);
    }
    std::vector<uint32_t> from, to;
    from.reserve(par->n_entities); to.reserve(par->n_entities);
    uint32_t missing = 0, cursor = 0;
    const uint32_t NOT_FOUND = 0xFFFFFFFFu;
    auto find_local = [&](uint32_t local) -> uint32_t {
        if (merge) {
            while (cursor < dst->n_entities && dst->h_ids[cursor] < local) cursor++;
            return (cursor < dst->n_entities && dst->h_ids[cursor] == local) ? cursor : NOT_FOUND;
        }
        auto it = where.find(local);
        return it == where.end() ? NOT_FOUND : it->second;
    };
    for (uint32_t i = 0; i < par->n_entities; i++) {
        const uint32_t id = par->h_ids[i], link = id & 0xFu;
        if (link >= ICX_NULL_CHAINLINK) continue;                                   // decrement_causal_chain_link fails: non-temporal, :213-219
        if (link == 0u)                                                             // ecs_types.h:148 throws std::range_error
            return cfail(dst, ICX_E_INVALID, "extract: entity with causal chain-link 0 in the parallel world");
        if ((id >> 28) == ICX_NULL_TIMESLICE) continue;                             // created during the parallel run, :223-229
        const uint32_t local = id - 1u;                                             // chain-link - 1
        const uint32_t j = find_local(local);
        if (j == NOT_FOUND) { missing++; continue; }                                // :241-245 registers it: entity lifecycle stays on the host
        if (par->active_mask[i] != dst->active_mask[j])
            return cfail(dst, ICX_E_STATE, "extract: collider activation differs; re-upload the destination");
        from.push_back(i); to.push_back(j);
    }
    const uint32_t cnt = (uint32_t)from.size();
    if (n_missing) *n_missing = missing;
    if (!cnt) return ICX_OK;
    int rc = gather(par, from, 1);
    if (rc != ICX_OK) { dst->error = par->error; return rc; }
    cudaSetDevice(dst->device);
    cudaStream_t st = dst->stream;
    CKW(dst, dst->st_index.ensure((size_t)cnt * 4)); CKW(dst, dst->st_T.ensure((size_t)cnt * sizeof(IcxTransform)));
    CKW(dst, dst->cl_ts.ensure(cnt));
    if (par->have_P) CKW(dst, dst->st_P.ensure((size_t)cnt * sizeof(IcxPhysics)));
    if (par->have_C) CKW(dst, dst->st_C.ensure((size_t)cnt * sizeof(IcxColliderComponent)));
    CKW(dst, cudaMemcpyAsync(dst->st_index.ptr, to.data(), (size_t)cnt * 4, cudaMemcpyHostToDevice, st));
    CKW(dst, xcopy(dst, dst->st_T.ptr, par, par->cl_T.ptr, (size_t)cnt * sizeof(IcxTransform)));
    if (par->have_P) CKW(dst, xcopy(dst, dst->st_P.ptr, par, par->cl_P.ptr, (size_t)cnt * sizeof(IcxPhysics)));
    if (par->have_C) CKW(dst, xcopy(dst, dst->st_C.ptr, par, par->cl_C.ptr, (size_t)cnt * sizeof(IcxColliderComponent)));
    CKW(dst, xcopy(dst, dst->cl_ts.ptr, par, par->cl_ts.ptr, cnt));
    Mirror M = mirror(dst);
    if (!dst->have_P) M.P = nullptr;
    // deserialize_entity_components(..., ComponentCategory::all), :252
    launch_import_kernel(dst->W, M, (const uint32_t*)dst->st_index.ptr, nullptr, (const IcxTransform*)dst->st_T.ptr,
                         par->have_P ? (const IcxPhysics*)dst->st_P.ptr : nullptr,
                         par->have_C ? (const IcxColliderComponent*)dst->st_C.ptr : nullptr,
                         nullptr, nullptr, nullptr, cnt, 0, dst->have_C ? 1 : 0, par->have_C ? ++dst->c_epoch : dst->c_epoch, st);
    launch_extract_ts(dst->W, (const uint32_t*)dst->st_index.ptr, (const uint8_t*)dst->cl_ts.ptr, cnt, st);
    CKW(dst, cudaStreamSynchronize(st));
    CKW(dst, cudaGetLastError());
    if (n_extracted) *n_extracted = cnt;
    return ICX_OK;
}

} // extern "C"
