// C ABI of include/icx.h: world lifetime, AoS import/export, the tick driver.
// No CPU fallback: every compute entry point needs a CUDA device.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <cstring>
#include <new>
#include "icx_host.h"

using namespace icx;

namespace {

#define CK(call) do { cudaError_t _e = (call); if (_e != cudaSuccess) { return fail_cuda(w, _e, #call); } } while (0)

int fail(IcxWorld* w, int code, const std::string& msg) { if (w) w->error = msg; return code; }
int fail_cuda(IcxWorld* w, cudaError_t e, const char* what)
{
    if (w) w->error = std::string(what) + ": " + cudaGetErrorString(e);
    return ICX_E_CUDA;
}

template <typename T> cudaError_t alloc(IcxWorld* w, T*& p, size_t count)
{
    DeviceBuffer b; cudaError_t e = b.ensure(std::max<size_t>(count, 1) * sizeof(T));
    if (e != cudaSuccess) return e;
    w->pool.push_back(b); p = static_cast<T*>(b.ptr);
    return cudaSuccess;
}

cudaError_t alloc_state(IcxWorld* w, EntityState& s, size_t n)
{
    cudaError_t e;
    if ((e = alloc(w, s.pos, n))) return e;  if ((e = alloc(w, s.rot, n))) return e;
    if ((e = alloc(w, s.vel, n))) return e;  if ((e = alloc(w, s.acc, n))) return e;
    if ((e = alloc(w, s.angvel, n))) return e;  if ((e = alloc(w, s.angacc, n))) return e;
    return alloc(w, s.ts, n);
}

uint32_t next_pow2(uint32_t v) { uint32_t p = 1; while (p < v) p <<= 1; return p; }

int allocate(IcxWorld* w)
{
    DeviceWorld& W = w->W;
    const size_t ne = w->cap_entities, np = w->cap_packets;
    CK(alloc_state(w, W.in, ne)); CK(alloc_state(w, W.out, ne));
    CK(alloc(w, W.prm0, ne)); CK(alloc(w, W.prm1, ne)); CK(alloc(w, W.locked_pos, ne)); CK(alloc(w, W.locked_rot, ne));
    CK(alloc(w, W.entity_id, ne)); CK(alloc(w, W.first_packet, ne)); CK(alloc(w, W.pos0, ne)); CK(alloc(w, W.c_ver, ne));
    CK(alloc(w, W.pk_entity, np)); CK(alloc(w, W.pk_lpos, np)); CK(alloc(w, W.pk_lrot, np)); CK(alloc(w, W.pk_lscale, np));
    CK(alloc(w, W.pk_mat0, np)); CK(alloc(w, W.pk_damping, np)); CK(alloc(w, W.pk_minparam, np));
    CK(alloc(w, W.pk_d, np * 6)); CK(alloc(w, W.pk_lo, np)); CK(alloc(w, W.pk_hi, np)); CK(alloc(w, W.pk_cell, np));
    W.n_buckets = std::max<uint32_t>(next_pow2((uint32_t)std::min<size_t>(2 * np, 1u << 24)), 1024u);     // capacity; resized per upload
    w->cap_buckets = W.n_buckets;
    CK(alloc(w, W.cell_count, (size_t)W.n_buckets + 4)); CK(alloc(w, W.cell_fill, (size_t)W.n_buckets));
    CK(alloc(w, W.scan_partial, 4096 + 8));
    CK(alloc(w, W.s_lo, np)); CK(alloc(w, W.s_hi, np));
    CK(alloc(w, W.large_list, W.max_large));
    CK(alloc(w, W.pairs, W.max_pairs)); CK(alloc(w, W.tight, W.max_pairs)); CK(alloc(w, W.push, ne)); CK(alloc(w, W.pushv, ne)); CK(alloc(w, W.static_pairs, W.max_static));
    CK(alloc(w, W.bm, ne)); CK(alloc(w, W.bm_keep, ne)); CK(alloc(w, W.bm_need, ne)); CK(alloc(w, W.mover_list, ne)); CK(alloc(w, W.xpairs, W.max_xpairs));
    CK(cudaMemset(W.bm_keep, 0, ne * sizeof(float))); CK(cudaMemset(W.bm_need, 0, ne * sizeof(uint32_t)));
    CK(alloc(w, W.dirty_slot, ne)); CK(alloc(w, W.dirty_list, ne));
    W.max_nbr = (uint32_t)std::min<size_t>(ne * (size_t)W.max_neighbors, 0xFFFF0000u);
    CK(alloc(w, W.nbr, (size_t)W.max_nbr)); CK(alloc(w, W.nbr_count, 2 * ne)); CK(alloc(w, W.nbr_fill, 2 * ne)); CK(alloc(w, W.slot_flag, ne)); CK(alloc(w, W.slot_info, ne));
    CK(alloc(w, W.head[0], ne)); CK(alloc(w, W.head[1], ne));
    CK(alloc(w, W.active[0], ne)); CK(alloc(w, W.active[1], ne)); CK(alloc(w, W.items, ne + W.max_static));
    CK(alloc(w, W.contacts, W.max_contacts)); CK(alloc(w, W.intercepts, W.max_contacts)); CK(alloc(w, W.intercept_key, W.max_contacts)); CK(alloc(w, W.grounds, W.max_contacts));
    CK(alloc(w, W.counters, CTR_TOTAL)); CK(alloc(w, W.flags, 4));
    CK(cudaMemset(W.counters, 0, CTR_TOTAL * sizeof(uint32_t))); CK(cudaMemset(W.flags, 0, 16));
    W.trace = nullptr;
    w->resolve_blocks = resolve_grid_blocks(w->device, 1); w->prepare_blocks = resolve_grid_blocks(w->device, 0);
    { int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, w->device); w->narrow_blocks = sms * 8; }
    if (w->resolve_blocks <= 0 || w->prepare_blocks <= 0) return fail(w, ICX_E_CUDA, "k_resolve cannot be made resident");
    CK(cudaMallocHost((void**)&w->h_counters, CTR_TOTAL * sizeof(uint32_t)));
    CK(cudaMallocHost((void**)&w->h_flags, 16));
    return ICX_OK;
}

// grow a single pooled array (used by the retry path)
template <typename T> int regrow(IcxWorld* w, T*& p, size_t count)
{
    for (auto& b : w->pool) if (b.ptr == p) { b.release(); CK(b.ensure(count * sizeof(T))); p = static_cast<T*>(b.ptr); return ICX_OK; }
    return fail(w, ICX_E_STATE, "regrow: buffer not in pool");
}

Mirror mirror_of(IcxWorld* w)
{
    Mirror m;
    m.ids = (uint32_t*)w->mir_ids.ptr; m.T = (IcxTransform*)w->mir_T.ptr; m.P = (IcxPhysics*)w->mir_P.ptr;
    m.C = (IcxColliderComponent*)w->mir_C.ptr; m.hasP = (uint8_t*)w->mir_hasP.ptr;
    return m;
}

void set_cell(IcxWorld* w, float cell) { w->W.cell_size = cell; w->W.inv_cell = 1.0f / cell; }

// One tick attempt on the stream: integrate+prep, grid, pairs, speculative narrowphase, ordered
// resolve of the dirty bodies.  One host sync per tick (the counters / retry flags).
int run_tick_once(IcxWorld* w, double dt, int do_euler, int do_collide, uint32_t* raised)
{
    DeviceWorld& W = w->W;
    cudaStream_t st = w->stream;
    CK(cudaMemsetAsync(W.counters, 0, CTR_TOTAL * sizeof(uint32_t), st));
    CK(cudaMemsetAsync(W.flags, 0, 16, st));
    CK(cudaMemsetAsync(W.cell_count, 0, ((size_t)W.n_buckets + 4) * sizeof(uint32_t), st));
    CK(cudaMemsetAsync(W.cell_fill, 0, (size_t)W.n_buckets * sizeof(uint32_t), st));
    const bool prof = w->profile && do_collide && W.n_packets;
    if (prof) cudaEventRecord(w->pev[0], st);
    launch_integrate_prep(W, dt, do_euler, st);
    w->launches += 1;
    if (prof) cudaEventRecord(w->pev[1], st);
    *raised = 0;
    if (do_collide && W.n_packets) {
        launch_grid(W, st);
        if (prof) cudaEventRecord(w->pev[2], st);
        launch_pairs(W, st);
        if (prof) cudaEventRecord(w->pev[3], st);
        launch_narrow(W, w->narrow_blocks, st);
        if (prof) cudaEventRecord(w->pev[4], st);
        CK(launch_resolve(W, dt, do_euler, w->prepare_blocks, w->resolve_blocks, st));
        w->launches += 8;
        if (prof) cudaEventRecord(w->pev[5], st);
    }
    CK(cudaMemcpyAsync(w->h_counters, W.counters, CTR_TOTAL * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(w->h_flags, W.flags, 16, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    CK(cudaGetLastError());
    *raised = w->h_flags[0];
    w->stats.n_rounds = w->h_counters[CTR_ROUNDS];
    if (prof) {
        for (int k = 0; k < ICX_N_PHASES; k++) { float ms = 0; cudaEventElapsedTime(&ms, w->pev[k], w->pev[k + 1]); w->phase_ms[k] += ms; }
        w->phase_ticks++;
    }
    return ICX_OK;
}

int run_tick(IcxWorld* w, double dt, int do_euler, int do_collide, bool advance_coherency)
{
    DeviceWorld& W = w->W;
    uint32_t last_raised = 0;
    for (int attempt = 0; attempt < 24; attempt++) {
        uint32_t raised = 0;
        int rc = run_tick_once(w, dt, do_euler, do_collide, &raised);
        if (rc != ICX_OK) return rc;
        if (!do_collide) raised &= ~(RF_CELL_TOO_SMALL | RF_LARGE_OVERFLOW);
        if (raised == 0) {
            std::swap(W.in, W.out);
            if (advance_coherency) W.coherency = (uint16_t)(W.coherency + 1);
            w->stats.ticks++;
            w->stats.n_large = std::min(w->h_counters[CTR_LARGE], W.max_large);
            w->stats.n_candidates = w->h_counters[CTR_CANDIDATES];
            w->stats.n_contacts = w->h_counters[CTR_CONTACTS];
            w->stats.n_interceptions = w->h_counters[CTR_INTERCEPTS];
            w->stats.n_dirty = w->h_counters[CTR_DIRTY]; w->stats.n_attempts = w->h_counters[CTR_ATTEMPTS] + 1;
            w->stats.n_movers = w->h_counters[CTR_MOVERS];
            w->totals[0] += 1; w->totals[1] += w->h_counters[CTR_CANDIDATES] + std::min(w->h_counters[CTR_XPAIRS], W.max_xpairs);
            w->totals[2] += w->h_counters[CTR_CONTACTS]; w->totals[3] += w->h_counters[CTR_ROUNDS];
            w->totals[4] += w->h_counters[CTR_ATTEMPTS] + 1; w->totals[5] += w->h_counters[CTR_DIRTY];
            w->stats.margin = W.margin; w->stats.cell_size = W.cell_size; w->stats.max_neighbors = W.max_neighbors;
            w->have_results = do_collide != 0;
            if (do_collide) return icx_mesh_pass(w);        // static triangle mesh (extension): after the pair loop, on the swapped state
            return ICX_OK;
        }
        // `in` is untouched: adjust the capacity / margin that was exceeded and run the tick again
        w->stats.n_retries++;
        last_raised = raised;
        if (getenv("ICX_DEBUG_RETRY"))
            fprintf(stderr, "[icx] tick %llu attempt %d: flags 0x%x dirty %u movers %u xpairs %u xmax %g violators %u candidates %u neighbors %u rounds %u\n",
                    (unsigned long long)w->stats.ticks, attempt, raised, w->h_counters[CTR_DIRTY], w->h_counters[CTR_MOVERS], w->h_counters[CTR_XPAIRS],
                    (double)*(float*)&w->h_counters[CTR_XMAX], w->h_counters[CTR_VIOLATORS], w->h_counters[CTR_CANDIDATES], W.max_neighbors, w->h_counters[CTR_ROUNDS]);
        if (raised & RF_CLIP_OVERFLOW) return fail(w, ICX_E_CAPACITY, "clipped manifold exceeded ICX_MAXV");
        if (raised & RF_NEIGHBOR_OVERFLOW) {
            if (W.max_neighbors >= 2048) return fail(w, ICX_E_CAPACITY, "candidate lists exceeded 2048 entries per entity on average (or 65535 in one list)");
            W.max_neighbors *= 2;
            W.max_nbr = (uint32_t)std::min<size_t>((size_t)w->cap_entities * W.max_neighbors, 0xFFFF0000u);
            rc = regrow(w, W.nbr, (size_t)W.max_nbr); if (rc) return rc;
        }
        if (raised & RF_PAIR_OVERFLOW) {
            if (w->h_counters[CTR_CANDIDATES] > W.max_pairs) {
                W.max_pairs = std::max<uint32_t>(W.max_pairs * 2, w->h_counters[CTR_CANDIDATES] + w->h_counters[CTR_CANDIDATES] / 4);
                rc = regrow(w, W.pairs, W.max_pairs); if (rc) return rc;
                rc = regrow(w, W.tight, W.max_pairs); if (rc) return rc;
            }
            if (w->h_counters[CTR_XPAIRS] > W.max_xpairs) {
                W.max_xpairs = std::max<uint32_t>(W.max_xpairs * 2, w->h_counters[CTR_XPAIRS] + w->h_counters[CTR_XPAIRS] / 4);
                rc = regrow(w, W.xpairs, W.max_xpairs); if (rc) return rc;
            }
            if (w->h_counters[CTR_STATIC] > W.max_static) {
                W.max_static = w->h_counters[CTR_STATIC] * 2;
                rc = regrow(w, W.static_pairs, W.max_static); if (rc) return rc;
                rc = regrow(w, W.items, (size_t)w->cap_entities + W.max_static); if (rc) return rc;     // static pairs ride in the first round's item list
            }
        }
        if (raised & RF_MARGIN_VIOLATED) return fail(w, ICX_E_CAPACITY, "resolve: per-body displacement budgets did not settle within the attempt limit");
        if (raised & RF_CELL_TOO_SMALL) {
            float need; std::memcpy(&need, &w->h_counters[CTR_NEED_CELL], 4);
            set_cell(w, std::max(need, W.cell_size) * 1.05f);
        }
        if (raised & RF_CONTACT_OVERFLOW) {
            W.max_contacts *= 2;
            rc = regrow(w, W.contacts, W.max_contacts); if (rc) return rc;
            rc = regrow(w, W.intercepts, W.max_contacts); if (rc) return rc;
            rc = regrow(w, W.intercept_key, W.max_contacts); if (rc) return rc;
            rc = regrow(w, W.grounds, W.max_contacts); if (rc) return rc;
        }
        if (raised & RF_LARGE_OVERFLOW) {
            W.max_large *= 2;
            rc = regrow(w, W.large_list, W.max_large); if (rc) return rc;
        }
        if (raised == RF_ROUNDS_EXCEEDED) return fail(w, ICX_E_CAPACITY, "resolve: dependency chain longer than ICX_MAX_ROUNDS");
    }
    { char msg[160]; snprintf(msg, sizeof msg, "tick did not converge after 24 capacity / budget retries (last flags 0x%x, %u bodies over budget)", last_raised, w->h_counters[CTR_VIOLATORS]);
      return fail(w, ICX_E_CAPACITY, msg); }
}

} // namespace

int icx_host_masks(IcxWorld* w)
{
    if (w->masks_on_host) return ICX_OK;
    w->active_mask.assign(w->n_entities, 0);
    if (w->n_entities) {
        cudaSetDevice(w->device);
        CK(cudaMemcpyAsync(w->active_mask.data(), w->d_mask.ptr, w->n_entities, cudaMemcpyDeviceToHost, w->stream));
        CK(cudaStreamSynchronize(w->stream));
    }
    w->masks_on_host = true;
    return ICX_OK;
}

extern "C" {

int icx_abi_version(void) { return ICX_ABI_VERSION; }

int icx_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

const char* icx_status_string(int status)
{
    switch (status) {
    case ICX_OK: return "ok";
    case ICX_E_INVALID: return "invalid argument";
    case ICX_E_CUDA: return "CUDA error";
    case ICX_E_CAPACITY: return "capacity exceeded";
    case ICX_E_NOMEM: return "out of memory";
    case ICX_E_STATE: return "wrong state for this call";
    default: return "unknown status";
    }
}

void icx_config_default(IcxConfig* cfg)
{
    if (!cfg) return;
    std::memset(cfg, 0, sizeof(*cfg));
    cfg->struct_size = sizeof(IcxConfig);
    cfg->max_entities = 1u << 16;
    cfg->host_timeslice = 0;
    cfg->gravity[0] = 0.0f; cfg->gravity[1] = -10.0f; cfg->gravity[2] = 0.0f;     // euler_physics_relay.h:32
}

int icx_world_create(const IcxConfig* cfg, IcxWorld** out)
{
    if (!cfg || !out || cfg->struct_size != sizeof(IcxConfig) || cfg->max_entities == 0 || cfg->max_entities > (1u << 24))      /* device scans are two-level: 4096 x 4096 */
        return ICX_E_INVALID;
    *out = nullptr;
    if (icx_device_count() <= cfg->device) return ICX_E_CUDA;      // no CPU fallback
    IcxWorld* w = new (std::nothrow) IcxWorld();
    if (!w) return ICX_E_NOMEM;
    w->cfg = *cfg; w->device = cfg->device;
    std::memset(&w->W, 0, sizeof(w->W)); std::memset(&w->stats, 0, sizeof(w->stats));
    int rc = ICX_OK;
    do {
        if (cudaSetDevice(w->device) != cudaSuccess) { rc = ICX_E_CUDA; break; }
        if (cudaStreamCreateWithFlags(&w->stream, cudaStreamNonBlocking) != cudaSuccess) { rc = ICX_E_CUDA; break; }
        cudaEventCreate(&w->ev0); cudaEventCreate(&w->ev1);
        DeviceWorld& W = w->W;
        w->cap_entities = cfg->max_entities; w->cap_packets = 2 * cfg->max_entities;
        W.margin = cfg->broadphase_margin > 0 ? cfg->broadphase_margin : 0.05f;
        set_cell(w, cfg->cell_size > 0 ? cfg->cell_size : 1.0f);
        W.max_neighbors = cfg->max_neighbors ? std::min<uint32_t>(cfg->max_neighbors, 2048) : 24;
        W.max_pairs = cfg->max_pairs ? cfg->max_pairs : std::max<uint32_t>(8 * cfg->max_entities, 4096);
        W.max_static = 4096;
        W.max_xpairs = std::max<uint32_t>(cfg->max_entities, 4096);
        W.max_large = cfg->max_large ? cfg->max_large : 1024;
        W.max_contacts = cfg->max_contacts ? cfg->max_contacts : std::max<uint32_t>(2 * cfg->max_entities, 1024);
        W.gx = (cfg->flags & ICX_F_NO_GRAVITY) ? 0.0f : cfg->gravity[0];
        W.gy = (cfg->flags & ICX_F_NO_GRAVITY) ? 0.0f : cfg->gravity[1];
        W.gz = (cfg->flags & ICX_F_NO_GRAVITY) ? 0.0f : cfg->gravity[2];
        W.host_timeslice = cfg->host_timeslice; W.cfg_flags = cfg->flags; W.worldline_stride = cfg->worldline_stride;
        rc = allocate(w);
    } while (0);
    if (rc != ICX_OK) { icx_world_destroy(w); return rc; }
    *out = w;
    return ICX_OK;
}

void icx_world_destroy(IcxWorld* w)
{
    if (!w) return;
    cudaSetDevice(w->device);
    if (w->stream) cudaStreamSynchronize(w->stream);
    for (auto& b : w->pool) b.release();
    DeviceBuffer* extra[] = {&w->mir_ids, &w->mir_T, &w->mir_P, &w->mir_C, &w->mir_hasP, &w->st_ts, &w->st_tc, &w->st_index, &w->st_T, &w->st_P, &w->st_C,
                             &w->d_mask, &w->st_scan, &w->cl_idx, &w->cl_ids, &w->cl_T, &w->cl_P, &w->cl_C, &w->cl_hasP, &w->cl_ts, &w->cl_tc,
                             &w->pj_proj, &w->pj_index, &w->pj_condemned, &w->pj_count,
                             &w->pend_rows, &w->pend_flag, &w->pend_list, &w->pend_count};
    for (auto* b : extra) b->release();
    icx_mesh_release(w);
    if (w->h_counters) cudaFreeHost(w->h_counters);
    if (w->h_flags) cudaFreeHost(w->h_flags);
    if (w->ev0) cudaEventDestroy(w->ev0);
    if (w->ev1) cudaEventDestroy(w->ev1);
    if (w->pj_ev0) cudaEventDestroy(w->pj_ev0);
    if (w->pj_ev1) cudaEventDestroy(w->pj_ev1);
    for (auto& e : w->pev) if (e) cudaEventDestroy(e);
    if (w->stream) cudaStreamDestroy(w->stream);
    delete w;
}

const char* icx_last_error(const IcxWorld* w) { return w ? w->error.c_str() : "null world"; }

int icx_device_share_set(IcxWorld* w, uint32_t parts)
{
    if (!w || parts == 0) return ICX_E_INVALID;
    const int full_r = resolve_grid_blocks(w->device, 1), full_p = resolve_grid_blocks(w->device, 0);
    if (full_r <= 0 || full_p <= 0) return fail(w, ICX_E_CUDA, "k_resolve cannot be made resident");
    w->resolve_blocks = std::max(1, full_r / (int)parts);
    w->prepare_blocks = std::max(1, full_p / (int)parts);
    return ICX_OK;
}

void* icx_host_alloc(size_t bytes)
{
    void* p = nullptr;
    if (icx_device_count() <= 0) return nullptr;
    if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocPortable) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    return p;
}
void icx_host_free(void* p) { if (p) cudaFreeHost(p); }

int icx_bodies_upload(IcxWorld* w, uint32_t n, const IcxEntity* entity, const IcxTransform* T, const IcxPhysics* P,
                      const uint8_t* has_physics, const IcxColliderComponent* C, const uint8_t* tstate, const uint16_t* tcoh)
{
    if (!w) return ICX_E_INVALID;
    if (!entity || !T || n > w->cap_entities) return fail(w, ICX_E_INVALID, "upload: null ids/transforms or n > max_entities");
    cudaSetDevice(w->device);
    DeviceWorld& W = w->W;
    cudaStream_t st = w->stream;
    w->h_ids.assign(entity, entity + n);
    w->masks_on_host = false;
    CK(w->d_mask.ensure(w->cap_entities)); CK(w->st_scan.ensure(((size_t)w->cap_entities + 4) * 4));
    CK(w->mir_ids.ensure((size_t)w->cap_entities * 4)); CK(w->mir_T.ensure((size_t)w->cap_entities * sizeof(IcxTransform)));
    CK(w->mir_hasP.ensure(w->cap_entities));
    if (P) CK(w->mir_P.ensure((size_t)w->cap_entities * sizeof(IcxPhysics)));
    if (C) CK(w->mir_C.ensure((size_t)w->cap_entities * sizeof(IcxColliderComponent)));
    CK(w->st_ts.ensure(w->cap_entities)); CK(w->st_tc.ensure((size_t)w->cap_entities * 2));
    w->have_P = P != nullptr; w->have_C = C != nullptr;
    CK(cudaMemcpyAsync(w->mir_ids.ptr, entity, (size_t)n * 4, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(w->mir_T.ptr, T, (size_t)n * sizeof(IcxTransform), cudaMemcpyHostToDevice, st));
    if (P) CK(cudaMemcpyAsync(w->mir_P.ptr, P, (size_t)n * sizeof(IcxPhysics), cudaMemcpyHostToDevice, st));
    if (C) CK(cudaMemcpyAsync(w->mir_C.ptr, C, (size_t)n * sizeof(IcxColliderComponent), cudaMemcpyHostToDevice, st));
    if (has_physics) CK(cudaMemcpyAsync(w->mir_hasP.ptr, has_physics, n, cudaMemcpyHostToDevice, st));
    if (tstate) CK(cudaMemcpyAsync(w->st_ts.ptr, tstate, n, cudaMemcpyHostToDevice, st));
    if (tcoh) CK(cudaMemcpyAsync(w->st_tc.ptr, tcoh, (size_t)n * 2, cudaMemcpyHostToDevice, st));
    // packet structure (one packet per active collider, submission order) on the device, from the uploaded components
    launch_packet_structure(C ? (const IcxColliderComponent*)w->mir_C.ptr : nullptr, n, (uint32_t*)w->st_scan.ptr, W.scan_partial,
                            (uint8_t*)w->d_mask.ptr, W.first_packet, st);
    CK(cudaMemsetAsync(W.counters, 0, CTR_TOTAL * sizeof(uint32_t), st));
    CK(cudaMemsetAsync(W.bm_keep, 0, (size_t)w->cap_entities * sizeof(float), st)); CK(cudaMemsetAsync(W.bm_need, 0, (size_t)w->cap_entities * sizeof(uint32_t), st));
    W.n_entities = n; w->n_entities = n;
    w->n_projectiles = 0; w->pj_updated = false;     // projectile rows index the previous upload order
    Mirror M = mirror_of(w);
    if (!P) M.P = nullptr;
    launch_import_kernel(W, M, nullptr, M.ids, M.T, M.P, C ? M.C : nullptr,
                         has_physics ? M.hasP : nullptr, tstate ? (const uint8_t*)w->st_ts.ptr : nullptr,
                         tcoh ? (const uint16_t*)w->st_tc.ptr : nullptr, n, 1, C ? 1 : 0, ++w->c_epoch, st);
    w->pend_any = false;
    CK(cudaMemcpyAsync(w->h_counters, W.counters, CTR_TOTAL * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(w->h_flags + 1, (uint32_t*)w->st_scan.ptr + n, 4, cudaMemcpyDeviceToHost, st));      // packet count
    CK(cudaStreamSynchronize(st));          // the caller's arrays may be released after return
    CK(cudaGetLastError());
    const uint32_t np = w->h_flags[1];
    W.n_packets = np; w->n_packets = np;
    if (np > w->cap_packets || np >= (1u << 28)) { W.n_entities = 0; W.n_packets = 0; w->n_entities = 0; w->n_packets = 0; return fail(w, ICX_E_CAPACITY, "upload: too many active colliders"); }
    icx_size_buckets(w);
    if (!(w->cfg.cell_size > 0)) {
        float ext; std::memcpy(&ext, &w->h_counters[CTR_EXTENT], 4);
        if (!(ext > 0)) ext = 1.0f;
        set_cell(w, (ext + 2.0f * W.margin) * 1.02f);
    }
    w->stats.n_entities = n; w->stats.n_packets = np;
    w->have_results = false;
    return ICX_OK;
}

int icx_bodies_download(IcxWorld* w, uint32_t n, IcxTransform* T, IcxPhysics* P, IcxColliderComponent* C, uint8_t* tstate, uint16_t* tcoh)
{
    if (!w) return ICX_E_INVALID;
    if (n != w->n_entities) return fail(w, ICX_E_INVALID, "download: n differs from the uploaded entity count");
    if ((P && !w->have_P) || (C && !w->have_C)) return fail(w, ICX_E_STATE, "download: component was never uploaded");
    cudaSetDevice(w->device);
    cudaStream_t st = w->stream;
    launch_export_kernel(w->W, mirror_of(w), T != nullptr, P != nullptr, C != nullptr,
                         tstate ? (uint8_t*)w->st_ts.ptr : nullptr, tcoh ? (uint16_t*)w->st_tc.ptr : nullptr, n, st);
    if (T) CK(cudaMemcpyAsync(T, w->mir_T.ptr, (size_t)n * sizeof(IcxTransform), cudaMemcpyDeviceToHost, st));
    if (P) CK(cudaMemcpyAsync(P, w->mir_P.ptr, (size_t)n * sizeof(IcxPhysics), cudaMemcpyDeviceToHost, st));
    if (C) CK(cudaMemcpyAsync(C, w->mir_C.ptr, (size_t)n * sizeof(IcxColliderComponent), cudaMemcpyDeviceToHost, st));
    if (tstate) CK(cudaMemcpyAsync(tstate, w->st_ts.ptr, n, cudaMemcpyDeviceToHost, st));
    if (tcoh) CK(cudaMemcpyAsync(tcoh, w->st_tc.ptr, (size_t)n * 2, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    CK(cudaGetLastError());
    return ICX_OK;
}

int icx_bodies_update(IcxWorld* w, uint32_t count, const uint32_t* index, const IcxTransform* T, const IcxPhysics* P, const IcxColliderComponent* C)
{
    if (!w) return ICX_E_INVALID;
    if (count > w->n_entities) return fail(w, ICX_E_INVALID, "update: more rows than uploaded entities");
    if ((P && !w->have_P) || (C && !w->have_C)) return fail(w, ICX_E_STATE, "update: component was never uploaded");
    if (!count) return ICX_OK;
    if (C) { int mrc = icx_host_masks(w); if (mrc != ICX_OK) return mrc; }
    if (index || C) for (uint32_t i = 0; i < count; i++) {
        const uint32_t e = index ? index[i] : i;
        if (e >= w->n_entities) return fail(w, ICX_E_INVALID, "update: index out of range");
        if (C) {
            uint8_t m = 0;
            for (int s = 0; s < ICX_COLLIDERS_PER_ENTITY; s++) if (C[i].colliders[s].isActive) m |= (uint8_t)(1u << s);
            if (m != w->active_mask[e]) return fail(w, ICX_E_STATE, "update: collider activation changed; call icx_bodies_upload");
        }
    }
    cudaSetDevice(w->device);
    cudaStream_t st = w->stream;
    if (index) {
        CK(w->st_index.ensure((size_t)count * 4));
        CK(cudaMemcpyAsync(w->st_index.ptr, index, (size_t)count * 4, cudaMemcpyHostToDevice, st));
    }
    if (T) { CK(w->st_T.ensure((size_t)count * sizeof(IcxTransform))); CK(cudaMemcpyAsync(w->st_T.ptr, T, (size_t)count * sizeof(IcxTransform), cudaMemcpyHostToDevice, st)); }
    if (P) { CK(w->st_P.ensure((size_t)count * sizeof(IcxPhysics))); CK(cudaMemcpyAsync(w->st_P.ptr, P, (size_t)count * sizeof(IcxPhysics), cudaMemcpyHostToDevice, st)); }
    if (C) { CK(w->st_C.ensure((size_t)count * sizeof(IcxColliderComponent))); CK(cudaMemcpyAsync(w->st_C.ptr, C, (size_t)count * sizeof(IcxColliderComponent), cudaMemcpyHostToDevice, st)); }
    Mirror M = mirror_of(w);
    if (!w->have_P) M.P = nullptr;
    launch_import_kernel(w->W, M, index ? (const uint32_t*)w->st_index.ptr : nullptr, nullptr, T ? (const IcxTransform*)w->st_T.ptr : nullptr,
                         P ? (const IcxPhysics*)w->st_P.ptr : nullptr, C ? (const IcxColliderComponent*)w->st_C.ptr : nullptr,
                         nullptr, nullptr, nullptr, count, 0, w->have_C ? 1 : 0, C ? ++w->c_epoch : w->c_epoch, st);
    CK(cudaStreamSynchronize(st));
    CK(cudaGetLastError());
    return ICX_OK;
}

int icx_motion_download(IcxWorld* w, uint32_t n, IcxMotionState* out)
{
    if (!w || (!out && n)) return ICX_E_INVALID;
    if (n != w->n_entities) return fail(w, ICX_E_INVALID, "motion_download: n differs from the uploaded entity count");
    if (!n) return ICX_OK;
    cudaSetDevice(w->device);
    cudaStream_t st = w->stream;
    CK(w->st_P.ensure((size_t)n * sizeof(IcxMotionState)));
    launch_motion_export(w->W, (IcxMotionState*)w->st_P.ptr, n, st);
    CK(cudaMemcpyAsync(out, w->st_P.ptr, (size_t)n * sizeof(IcxMotionState), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    CK(cudaGetLastError());
    return ICX_OK;
}

void icx_wire_layout_default(IcxWireLayout* l)
{
    if (!l) return;
    l->event_id = 102; l->type_transform = 0; l->type_physics = 6; l->type_collider = 7; l->increment_link = 1;     // event_types.h:102, hard_config_cosmos.h:47-58
}

int icx_entity_update_messages(IcxWorld* w, const IcxWireLayout* layout, uint16_t coherency, void* out, size_t capacity, size_t* stride_bytes, uint32_t* n_messages)
{
    if (!w || !layout || !out) return ICX_E_INVALID;
    const uint32_t n = w->n_entities;
    if (stride_bytes) *stride_bytes = ICX_WIRE_MAX_BYTES;
    if (n_messages) *n_messages = n;
    if ((size_t)n * ICX_WIRE_MAX_BYTES > capacity) return fail(w, ICX_E_CAPACITY, "entity_update_messages: buffer too small");
    if (layout->type_transform > 31 || layout->type_physics > 31 || layout->type_collider > 31) return fail(w, ICX_E_INVALID, "entity_update_messages: component type > 31");
    for (uint32_t i = 0; i < n; i++) if (((w->h_ids[i] >> 4) & 0xFFFFFFu) > 0xFFu) return fail(w, ICX_E_INVALID, "entity_update_messages: genesis id does not fit the reference's 16-bit Entity");
    if (!n) return ICX_OK;
    cudaSetDevice(w->device);
    cudaStream_t st = w->stream;
    CK(w->st_C.ensure((size_t)n * ICX_WIRE_MAX_BYTES));
    launch_export_kernel(w->W, mirror_of(w), 1, w->have_P ? 1 : 0, w->have_C ? 1 : 0, nullptr, nullptr, n, st);
    Mirror M = mirror_of(w);
    launch_wire_messages(w->W, M, *layout, coherency, w->have_C ? 1 : 0, (uint32_t*)w->st_C.ptr, n, st);
    CK(cudaMemcpyAsync(out, w->st_C.ptr, (size_t)n * ICX_WIRE_MAX_BYTES, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    CK(cudaGetLastError());
    return ICX_OK;
}

int icx_coherency_set(IcxWorld* w, uint16_t c) { if (!w) return ICX_E_INVALID; w->W.coherency = c; return ICX_OK; }
int icx_coherency_get(IcxWorld* w, uint16_t* c) { if (!w || !c) return ICX_E_INVALID; *c = w->W.coherency; return ICX_OK; }

static int timed_ticks(IcxWorld* w, double dt, uint32_t n_ticks, int do_euler, int do_collide, bool advance)
{
    if (!w) return ICX_E_INVALID;
    cudaSetDevice(w->device);
    cudaEventRecord(w->ev0, w->stream);
    for (uint32_t t = 0; t < n_ticks; t++) { int rc = run_tick(w, dt, do_euler, do_collide, advance); if (rc != ICX_OK) return rc; }
    cudaEventRecord(w->ev1, w->stream);
    cudaEventSynchronize(w->ev1);
    cudaEventElapsedTime(&w->last_ms, w->ev0, w->ev1);
    return ICX_OK;
}

int icx_step(IcxWorld* w, double dt) { return timed_ticks(w, dt, 1, 1, 1, true); }
int icx_step_n(IcxWorld* w, double dt, uint32_t n_ticks) { return timed_ticks(w, dt, n_ticks, 1, 1, true); }
int icx_euler_engage(IcxWorld* w, double dt) { return timed_ticks(w, dt, 1, 1, 0, false); }
int icx_collision_engage(IcxWorld* w) { return timed_ticks(w, 0.0, 1, 0, 1, false); }

int icx_sync(IcxWorld* w)
{
    if (!w) return ICX_E_INVALID;
    cudaSetDevice(w->device);
    CK(cudaStreamSynchronize(w->stream));
    CK(cudaGetLastError());
    return ICX_OK;
}

int icx_last_step_ms(IcxWorld* w, float* ms) { if (!w || !ms) return ICX_E_INVALID; *ms = w->last_ms; return ICX_OK; }
void* icx_world_stream(IcxWorld* w) { return w ? (void*)w->stream : nullptr; }

int icx_profile_enable(IcxWorld* w, int on)
{
    if (!w) return ICX_E_INVALID;
    cudaSetDevice(w->device);
    if (on && !w->pev[0]) for (auto& e : w->pev) cudaEventCreate(&e);
    w->profile = on != 0;
    for (auto& m : w->phase_ms) m = 0;
    for (auto& t : w->totals) t = 0;
    w->phase_ticks = 0;
    return ICX_OK;
}

int icx_phase_ms(IcxWorld* w, double* ms, uint64_t* n_ticks, uint64_t* n_launches)
{
    if (!w) return ICX_E_INVALID;
    if (ms) for (int k = 0; k < ICX_N_PHASES; k++) ms[k] = w->phase_ms[k];
    if (n_ticks) *n_ticks = w->phase_ticks;
    if (n_launches) *n_launches = w->launches;
    return ICX_OK;
}

int icx_totals_fetch(IcxWorld* w, uint64_t* out6)
{
    if (!w || !out6) return ICX_E_INVALID;
    for (int k = 0; k < 6; k++) out6[k] = w->totals[k];
    return ICX_OK;
}

int icx_debug_trace(IcxWorld* w, int enable, uint64_t* out, size_t cap, size_t* n)
{
    if (!w) return ICX_E_INVALID;
    cudaSetDevice(w->device);
    if (enable && !w->W.trace) { CK(alloc(w, w->W.trace, ICX_TRACE_CAP)); CK(cudaMemset(w->W.trace, 0, ICX_TRACE_CAP * 8)); }
    if (n) *n = 0;
    if (out && w->W.trace) {
        std::vector<unsigned long long> tmp(ICX_TRACE_CAP);
        CK(cudaMemcpyAsync(tmp.data(), w->W.trace, ICX_TRACE_CAP * 8, cudaMemcpyDeviceToHost, w->stream));
        CK(cudaStreamSynchronize(w->stream));
        const size_t used = std::min<size_t>(tmp[0], ICX_TRACE_CAP);
        for (size_t i = 2; i < used && i - 2 < cap; i++) out[i - 2] = tmp[i];
        if (n) *n = used >= 2 ? used - 2 : 0;
    }
    if (!enable) w->W.trace = nullptr;       // stays allocated in the pool
    return ICX_OK;
}

int icx_debug_counters(IcxWorld* w, uint32_t* out32)
{
    if (!w || !out32) return ICX_E_INVALID;
    std::memcpy(out32, w->h_counters, CTR_TOTAL * sizeof(uint32_t));
    return ICX_OK;
}

int icx_stats_fetch(IcxWorld* w, IcxStats* out) { if (!w || !out) return ICX_E_INVALID; *out = w->stats; return ICX_OK; }

int icx_contacts_fetch(IcxWorld* w, IcxContact* out, size_t cap, size_t* n)
{
    if (!w || !n) return ICX_E_INVALID;
    if (!w->have_results) { *n = 0; return ICX_OK; }
    const size_t total = w->stats.n_contacts;
    *n = total;
    if (!out || !total) return ICX_OK;
    cudaSetDevice(w->device);
    std::vector<IcxContact> tmp(total);
    CK(cudaMemcpyAsync(tmp.data(), w->W.contacts, total * sizeof(IcxContact), cudaMemcpyDeviceToHost, w->stream));
    CK(cudaStreamSynchronize(w->stream));
    // the rounds emit hits out of order; the reference's order is lexicographic on the packet pair
    std::sort(tmp.begin(), tmp.end(), [](const IcxContact& a, const IcxContact& b) {
        return a.packetA < b.packetA || (a.packetA == b.packetA && a.packetB < b.packetB); });
    std::memcpy(out, tmp.data(), std::min(cap, total) * sizeof(IcxContact));
    return ICX_OK;
}

int icx_interceptions_fetch(IcxWorld* w, IcxInterception* out, size_t cap, size_t* n)
{
    if (!w || !n) return ICX_E_INVALID;
    if (!w->have_results) { *n = 0; return ICX_OK; }
    const size_t total = w->stats.n_interceptions;
    *n = total;
    if (!out || !total) return ICX_OK;
    cudaSetDevice(w->device);
    std::vector<IcxInterception> ev(total); std::vector<uint2> key(total);
    CK(cudaMemcpyAsync(ev.data(), w->W.intercepts, total * sizeof(IcxInterception), cudaMemcpyDeviceToHost, w->stream));
    CK(cudaMemcpyAsync(key.data(), w->W.intercept_key, total * sizeof(uint2), cudaMemcpyDeviceToHost, w->stream));
    CK(cudaStreamSynchronize(w->stream));
    std::vector<uint32_t> order(total);
    for (size_t i = 0; i < total; i++) order[i] = (uint32_t)i;
    std::sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) {
        return key[a].x < key[b].x || (key[a].x == key[b].x && key[a].y < key[b].y); });
    for (size_t i = 0; i < std::min(cap, total); i++) out[i] = ev[order[i]];
    return ICX_OK;
}

int icx_grounds_fetch(IcxWorld* w, IcxGround* out, size_t cap, size_t* n)
{
    if (!w || !n) return ICX_E_INVALID;
    *n = 0;
    if (!w->have_results) return ICX_OK;
    const size_t total = std::min<size_t>(w->h_counters[CTR_GROUNDS], w->W.max_contacts);
    if (!total) return ICX_OK;
    cudaSetDevice(w->device);
    std::vector<IcxGround> tmp(total);
    CK(cudaMemcpyAsync(tmp.data(), w->W.grounds, total * sizeof(IcxGround), cudaMemcpyDeviceToHost, w->stream));
    CK(cudaStreamSynchronize(w->stream));
    // calls arrive out of order; the component keeps the LAST call in pair order (A's call of a pair precedes B's, relay :114-155)
    std::sort(tmp.begin(), tmp.end(), [](const IcxGround& a, const IcxGround& b) {
        if (a.entity != b.entity) return a.entity < b.entity;
        if (a.packetA != b.packetA) return a.packetA < b.packetA;
        if (a.packetB != b.packetB) return a.packetB < b.packetB;
        return a.side < b.side; });
    size_t k = 0;
    for (size_t i = 0; i < total; i++) {
        if (i + 1 < total && tmp[i + 1].entity == tmp[i].entity) continue;
        if (out && k < cap) out[k] = tmp[i];
        k++;
    }
    *n = k;
    return ICX_OK;
}

int icx_pairs_fetch(IcxWorld* w, IcxPair* out, size_t cap, size_t* n)
{
    if (!w || !n) return ICX_E_INVALID;
    if (!w->have_results) { *n = 0; return ICX_OK; }
    cudaSetDevice(w->device);
    const size_t total = w->stats.n_candidates;
    *n = total;
    if (!out || !total) return ICX_OK;
    std::vector<uint2> raw(total);
    CK(cudaMemcpyAsync(raw.data(), w->W.pairs, total * sizeof(uint2), cudaMemcpyDeviceToHost, w->stream));
    CK(cudaStreamSynchronize(w->stream));
    // the generator emits pairs in block order; canonical order is lexicographic on (a, b)
    std::vector<uint64_t> keys(total);
    for (size_t i = 0; i < total; i++) keys[i] = ((uint64_t)(raw[i].x & ICX_NB_INDEX_MASK) << 32) | (raw[i].y & ICX_NB_INDEX_MASK);
    std::sort(keys.begin(), keys.end());
    for (size_t i = 0; i < std::min(cap, total); i++) { out[i].a = (uint32_t)(keys[i] >> 32); out[i].b = (uint32_t)keys[i]; }
    return ICX_OK;
}

// The movers' extra candidates of the last tick (pairs only their displacement budgets create), canonical order.
int icx_extra_pairs_fetch(IcxWorld* w, IcxPair* out, size_t cap, size_t* n)
{
    if (!w || !n) return ICX_E_INVALID;
    if (!w->have_results) { *n = 0; return ICX_OK; }
    cudaSetDevice(w->device);
    const size_t total = std::min<size_t>(w->h_counters[CTR_XPAIRS], w->W.max_xpairs);
    std::vector<uint2> raw(total);
    if (total) {
        CK(cudaMemcpyAsync(raw.data(), w->W.xpairs, total * sizeof(uint2), cudaMemcpyDeviceToHost, w->stream));
        CK(cudaStreamSynchronize(w->stream));
    }
    std::vector<uint64_t> keys(total);
    for (size_t i = 0; i < total; i++) keys[i] = ((uint64_t)(raw[i].x & ICX_NB_INDEX_MASK) << 32) | (raw[i].y & ICX_NB_INDEX_MASK);
    std::sort(keys.begin(), keys.end());
    keys.erase(std::unique(keys.begin(), keys.end()), keys.end());       // rows that collide in the bucket hash are walked twice
    *n = keys.size();
    if (out) for (size_t i = 0; i < std::min(cap, keys.size()); i++) { out[i].a = (uint32_t)(keys[i] >> 32); out[i].b = (uint32_t)keys[i]; }
    return ICX_OK;
}

// Diagnostic: which = 0 predicted push, 1 budget, 2 largest displacement component static resolution produced in the last tick.
int icx_debug_array(IcxWorld* w, int which, float* out, size_t cap)
{
    if (!w || !out || which < 0 || which > 2) return ICX_E_INVALID;
    cudaSetDevice(w->device);
    const size_t n = std::min<size_t>(cap, w->n_entities);
    if (!n) return ICX_OK;
    if (which < 2) { CK(cudaMemcpyAsync(out, which ? w->W.bm : w->W.push, n * sizeof(float), cudaMemcpyDeviceToHost, w->stream)); CK(cudaStreamSynchronize(w->stream)); return ICX_OK; }
    std::vector<float4> a(n), b(n);
    CK(cudaMemcpyAsync(a.data(), w->W.in.pos, n * sizeof(float4), cudaMemcpyDeviceToHost, w->stream));
    CK(cudaMemcpyAsync(b.data(), w->W.pos0, n * sizeof(float4), cudaMemcpyDeviceToHost, w->stream));
    CK(cudaStreamSynchronize(w->stream));
    for (size_t i = 0; i < n; i++) out[i] = std::max(std::max(std::fabs(a[i].x - b[i].x), std::fabs(a[i].y - b[i].y)), std::fabs(a[i].z - b[i].z));
    return ICX_OK;
}

// Per-entity displacement budgets of the last tick (>= IcxStats.margin).
int icx_budgets_fetch(IcxWorld* w, float* out, size_t cap)
{
    if (!w || !out) return ICX_E_INVALID;
    cudaSetDevice(w->device);
    const size_t n = std::min<size_t>(cap, w->n_entities);
    if (n) { CK(cudaMemcpyAsync(out, w->W.bm, n * sizeof(float), cudaMemcpyDeviceToHost, w->stream)); CK(cudaStreamSynchronize(w->stream)); }
    return ICX_OK;
}

} // extern "C"
