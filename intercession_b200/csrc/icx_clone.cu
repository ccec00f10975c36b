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

// ---- entity lifecycle on the device (Cosmos::register_entity / condemn_entity as the timeline's ENTITY_CREATED / ENTITY_REMOVED
// handlers call them, server_network_dynamo.cpp:422-463; arrivals of time travellers, :748-812) ----
// Rows [n_old, n_old + n_app) of the AoS mirror and of in.ts hold the entities to add.  order[i] = the row (old or appended)
// that becomes row i of the new submission order.  The rows are gathered on the device into that order and re-imported like a
// fresh submission: existing entities never cross PCIe, and everything derived from the row order (packets, grid, candidate
// lists, collider-row versions) is rebuilt exactly as icx_bodies_upload of the merged arrays would build it.
int restructure(IcxWorld* w, const std::vector<uint32_t>& order)
{
    const uint32_t n_new = (uint32_t)order.size();
    std::vector<uint32_t> ids(n_new);           // (the caller put the ids of appended rows behind the live ones in h_ids)
    for (uint32_t i = 0; i < n_new; i++) ids[i] = w->h_ids[order[i]];
    int rc = gather(w, order, 1);               // refreshes the mirror rows of the old entities first; state words travel unchanged
    if (rc != ICX_OK) return rc;
    cudaSetDevice(w->device);
    cudaStream_t st = w->stream;
    DeviceWorld& W = w->W;
    const size_t cap = w->cap_entities;
    CKW(w, w->st_ts.ensure(cap)); CKW(w, w->st_tc.ensure(cap * 2));
    CKW(w, w->d_mask.ensure(cap)); CKW(w, w->st_scan.ensure((cap + 4) * 4));
    if (n_new) {
        CKW(w, cudaMemcpyAsync(w->mir_ids.ptr, w->cl_ids.ptr, (size_t)n_new * 4, cudaMemcpyDeviceToDevice, st));
        CKW(w, cudaMemcpyAsync(w->mir_T.ptr, w->cl_T.ptr, (size_t)n_new * sizeof(IcxTransform), cudaMemcpyDeviceToDevice, st));
        CKW(w, cudaMemcpyAsync(w->mir_hasP.ptr, w->cl_hasP.ptr, n_new, cudaMemcpyDeviceToDevice, st));
        if (w->have_P) CKW(w, cudaMemcpyAsync(w->mir_P.ptr, w->cl_P.ptr, (size_t)n_new * sizeof(IcxPhysics), cudaMemcpyDeviceToDevice, st));
        if (w->have_C) CKW(w, cudaMemcpyAsync(w->mir_C.ptr, w->cl_C.ptr, (size_t)n_new * sizeof(IcxColliderComponent), cudaMemcpyDeviceToDevice, st));
        CKW(w, cudaMemcpyAsync(w->st_ts.ptr, w->cl_ts.ptr, n_new, cudaMemcpyDeviceToDevice, st));
        CKW(w, cudaMemcpyAsync(w->st_tc.ptr, w->cl_tc.ptr, (size_t)n_new * 2, cudaMemcpyDeviceToDevice, st));
    }
    launch_packet_structure(w->have_C ? (const IcxColliderComponent*)w->mir_C.ptr : nullptr, n_new, (uint32_t*)w->st_scan.ptr, W.scan_partial,
                            (uint8_t*)w->d_mask.ptr, W.first_packet, st);
    CKW(w, cudaMemsetAsync(W.counters, 0, CTR_TOTAL * sizeof(uint32_t), st));
    CKW(w, cudaMemsetAsync(W.bm_keep, 0, cap * sizeof(float), st)); CKW(w, cudaMemsetAsync(W.bm_need, 0, cap * sizeof(uint32_t), st));
    W.n_entities = n_new; w->n_entities = n_new;
    w->n_projectiles = 0; w->pj_updated = false;
    w->h_ids = ids; w->masks_on_host = false;
    Mirror M = mirror(w);
    if (!w->have_P) M.P = nullptr;
    launch_import_kernel(W, M, nullptr, M.ids, M.T, M.P, w->have_C ? M.C : nullptr, M.hasP, (const uint8_t*)w->st_ts.ptr,
                         (const uint16_t*)w->st_tc.ptr, n_new, 1, w->have_C ? 1 : 0, ++w->c_epoch, st);
    w->pend_any = false;
    CKW(w, cudaMemcpyAsync(w->h_counters, W.counters, CTR_TOTAL * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    CKW(w, cudaMemcpyAsync(w->h_flags + 1, (uint32_t*)w->st_scan.ptr + n_new, 4, cudaMemcpyDeviceToHost, st));
    CKW(w, cudaStreamSynchronize(st));
    CKW(w, cudaGetLastError());
    const uint32_t np = n_new ? w->h_flags[1] : 0u;
    if (np > w->cap_packets || np >= (1u << 28)) return cfail(w, ICX_E_CAPACITY, "entity lifecycle: too many active colliders");
    W.n_packets = np; w->n_packets = np;
    icx_size_buckets(w);
    if (!(w->cfg.cell_size > 0)) {                                                  // same rule as icx_bodies_upload
        float ext; std::memcpy(&ext, &w->h_counters[CTR_EXTENT], 4);
        if (!(ext > 0)) ext = 1.0f;
        W.cell_size = (ext + 2.0f * W.margin) * 1.02f; W.inv_cell = 1.0f / W.cell_size;
    }
    w->stats.n_entities = n_new; w->stats.n_packets = np;
    w->have_results = false;
    return ICX_OK;
}

} // namespace

extern "C" {

int icx_entities_create(IcxWorld* w, uint32_t count, const IcxEntity* entity, const IcxTransform* T, const IcxPhysics* P,
                        const uint8_t* has_physics, const IcxColliderComponent* C, const uint8_t* tstate, const uint16_t* tcoh,
                        uint32_t* n_created)
{
    if (!w) return ICX_E_INVALID;
    if (n_created) *n_created = 0;
    if (!count) return ICX_OK;
    if (!entity || !T) return cfail(w, ICX_E_INVALID, "create: null ids / transforms");
    if ((w->have_P && !P) || (w->have_C && !C)) return cfail(w, ICX_E_INVALID, "create: the world holds physics / collider components, the new rows need them too");
    const uint32_t n_old = w->n_entities;
    // an id that exists already is left alone (server_network_dynamo.cpp:431-437 "excessive creation ... ignoring")
    const bool live_sorted = std::is_sorted(w->h_ids.begin(), w->h_ids.begin() + n_old);
    std::vector<uint32_t> take;                                                     // rows of the caller's arrays that are created
    if (live_sorted) {                                                              // (the usual case: one binary search per new id)
        std::vector<uint32_t> seen;
        for (uint32_t i = 0; i < count; i++) {
            if (std::binary_search(w->h_ids.begin(), w->h_ids.begin() + n_old, entity[i])) continue;
            if (std::find(seen.begin(), seen.end(), entity[i]) != seen.end()) continue;
            seen.push_back(entity[i]); take.push_back(i);
        }
    } else {
        std::unordered_map<uint32_t, uint32_t> have; have.reserve((size_t)n_old * 2 + count);
        for (uint32_t j = 0; j < n_old; j++) have.emplace(w->h_ids[j], j);
        for (uint32_t i = 0; i < count; i++) if (have.emplace(entity[i], n_old + (uint32_t)take.size()).second) take.push_back(i);
    }
    const uint32_t n_app = (uint32_t)take.size();
    if (!n_app) return ICX_OK;
    if ((size_t)n_old + n_app > w->cap_entities) return cfail(w, ICX_E_CAPACITY, "create: max_entities too small");
    cudaSetDevice(w->device);
    cudaStream_t st = w->stream;
    const size_t cap = w->cap_entities;
    CKW(w, w->mir_ids.ensure(cap * 4)); CKW(w, w->mir_T.ensure(cap * sizeof(IcxTransform))); CKW(w, w->mir_hasP.ensure(cap));
    if (w->have_P) CKW(w, w->mir_P.ensure(cap * sizeof(IcxPhysics)));
    if (w->have_C) CKW(w, w->mir_C.ensure(cap * sizeof(IcxColliderComponent)));
    std::vector<uint32_t> ts(n_app); std::vector<uint8_t> hp(n_app);
    w->h_ids.resize(n_old);
    for (uint32_t a = 0; a < n_app; a++) {                                          // appended behind the live rows (mirror and state word)
        const uint32_t i = take[a];
        w->h_ids.push_back(entity[i]);
        ts[a] = (uint32_t)(tstate ? tstate[i] : ICX_TS_MERGED) | ((uint32_t)(tcoh ? tcoh[i] : 0) << 16);
        hp[a] = (uint8_t)((has_physics ? has_physics[i] != 0 : P != nullptr) ? 1 : 0);
        CKW(w, cudaMemcpyAsync((uint32_t*)w->mir_ids.ptr + n_old + a, entity + i, 4, cudaMemcpyHostToDevice, st));
        CKW(w, cudaMemcpyAsync((IcxTransform*)w->mir_T.ptr + n_old + a, T + i, sizeof(IcxTransform), cudaMemcpyHostToDevice, st));
        if (w->have_P) CKW(w, cudaMemcpyAsync((IcxPhysics*)w->mir_P.ptr + n_old + a, P + i, sizeof(IcxPhysics), cudaMemcpyHostToDevice, st));
        if (w->have_C) CKW(w, cudaMemcpyAsync((IcxColliderComponent*)w->mir_C.ptr + n_old + a, C + i, sizeof(IcxColliderComponent), cudaMemcpyHostToDevice, st));
    }
    CKW(w, cudaMemcpyAsync((uint8_t*)w->mir_hasP.ptr + n_old, hp.data(), n_app, cudaMemcpyHostToDevice, st));
    CKW(w, cudaMemcpyAsync(w->W.in.ts + n_old, ts.data(), (size_t)n_app * 4, cudaMemcpyHostToDevice, st));
    CKW(w, cudaStreamSynchronize(st));                                              // ts / hp are locals
    // new submission order: ascending ids when the live ones are (std::set<Entity>, ecs/i_synchro.h:48), else appended
    std::vector<uint32_t> order(n_old + n_app);
    if (live_sorted) {                                                              // merge: the few new rows (sorted among themselves) into the live run
        std::vector<uint32_t> fresh(n_app);
        for (uint32_t a = 0; a < n_app; a++) fresh[a] = n_old + a;
        std::sort(fresh.begin(), fresh.end(), [&](uint32_t a, uint32_t b) { return w->h_ids[a] < w->h_ids[b]; });
        uint32_t i = 0, a = 0, o = 0;
        while (i < n_old || a < n_app) {
            if (a == n_app || (i < n_old && w->h_ids[i] < w->h_ids[fresh[a]])) order[o++] = i++;
            else order[o++] = fresh[a++];
        }
    } else for (uint32_t i = 0; i < n_old + n_app; i++) order[i] = i;
    const int rc = restructure(w, order);
    if (rc == ICX_OK && n_created) *n_created = n_app;
    return rc;
}

int icx_entity_ids_fetch(IcxWorld* w, IcxEntity* out, size_t cap, size_t* n_out)
{
    if (!w) return ICX_E_INVALID;
    if (n_out) *n_out = w->n_entities;
    if (out) std::memcpy(out, w->h_ids.data(), std::min<size_t>(cap, w->n_entities) * sizeof(IcxEntity));
    return ICX_OK;
}

int icx_entities_remove(IcxWorld* w, uint32_t count, const IcxEntity* entity, uint32_t* n_removed)
{
    if (!w) return ICX_E_INVALID;
    if (n_removed) *n_removed = 0;
    if (!count) return ICX_OK;
    if (!entity) return cfail(w, ICX_E_INVALID, "remove: null ids");
    std::vector<uint32_t> gone(entity, entity + count);
    std::sort(gone.begin(), gone.end());
    std::vector<uint32_t> order; order.reserve(w->n_entities);
    for (uint32_t j = 0; j < w->n_entities; j++)
        if (!std::binary_search(gone.begin(), gone.end(), w->h_ids[j])) order.push_back(j);
    const uint32_t removed = w->n_entities - (uint32_t)order.size();
    if (!removed) return ICX_OK;                                                    // condemning an entity that is not there does nothing
    const int rc = restructure(w, order);
    if (rc == ICX_OK && n_removed) *n_removed = removed;
    return rc;
}

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
    return icx_world_extract_ex(par, dst, 0u, n_extracted, n_missing);
}

int icx_world_extract_ex(IcxWorld* par, IcxWorld* dst, uint32_t flags, uint32_t* n_extracted, uint32_t* n_missing)
{
    if (!par || !dst || par == dst) return ICX_E_INVALID;
    if (flags & ICX_EXTRACT_REGISTER) {
        // :241-245 "entity with non-null host from past does not already exist in future, registering it": the absent ones are
        // created in dst first, from the parallel world's rows, {merged, 0} like Cosmos::register_entity leaves them; the loop
        // below then finds every id
        std::vector<uint32_t> have(dst->h_ids.begin(), dst->h_ids.begin() + dst->n_entities);
        std::sort(have.begin(), have.end());
        std::vector<uint32_t> rows;
        for (uint32_t i = 0; i < par->n_entities; i++) {
            const uint32_t id = par->h_ids[i], link = id & 0xFu;
            if (link >= ICX_NULL_CHAINLINK || link == 0u || (id >> 28) == ICX_NULL_TIMESLICE) continue;
            if (!std::binary_search(have.begin(), have.end(), id - 1u)) rows.push_back(i);
        }
        if (!rows.empty()) {
            const size_t k = rows.size();
            int rc = gather(par, rows, 1);
            if (rc != ICX_OK) { dst->error = par->error; return rc; }
            std::vector<uint32_t> ids(k); std::vector<IcxTransform> T(k); std::vector<IcxPhysics> P(k); std::vector<IcxColliderComponent> C(k);
            std::vector<uint8_t> hp(k), ts(k, ICX_TS_MERGED); std::vector<uint16_t> tc(k, 0);
            cudaSetDevice(par->device);
            CKW(par, cudaMemcpy(ids.data(), par->cl_ids.ptr, k * 4, cudaMemcpyDeviceToHost));
            CKW(par, cudaMemcpy(T.data(), par->cl_T.ptr, k * sizeof(IcxTransform), cudaMemcpyDeviceToHost));
            CKW(par, cudaMemcpy(hp.data(), par->cl_hasP.ptr, k, cudaMemcpyDeviceToHost));
            if (par->have_P) CKW(par, cudaMemcpy(P.data(), par->cl_P.ptr, k * sizeof(IcxPhysics), cudaMemcpyDeviceToHost));
            if (par->have_C) CKW(par, cudaMemcpy(C.data(), par->cl_C.ptr, k * sizeof(IcxColliderComponent), cudaMemcpyDeviceToHost));
            for (size_t i = 0; i < k; i++) ids[i] -= 1u;                            // chain-link - 1
            uint32_t made = 0;
            rc = icx_entities_create(dst, (uint32_t)k, ids.data(), T.data(), par->have_P ? P.data() : nullptr, hp.data(),
                                     par->have_C ? C.data() : nullptr, ts.data(), tc.data(), &made);
            if (rc != ICX_OK) return rc;
        }
    }
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
