// Host-side world object behind the C ABI (include/icx.h).  Not part of the ABI.
#pragma once
#include <cuda_runtime.h>
#include <string>
#include <vector>
#include "icx_world.cuh"

namespace icx {

// launchers implemented in the .cu translation units
void launch_integrate_prep(const DeviceWorld& W, double dt, int do_euler, cudaStream_t st);
void launch_grid(const DeviceWorld& W, cudaStream_t st);      // 4 kernels: scan x3 + scatter
void launch_pairs(const DeviceWorld& W, cudaStream_t st);
void launch_narrow(const DeviceWorld& W, int blocks, cudaStream_t st);
int resolve_grid_blocks(int device, int phase);               // co-resident blocks of the persistent resolve kernels (0 = prepare, 1 = rounds)
void launch_mesh_pass(const DeviceWorld& W, const DeviceMesh& M, int sms, cudaStream_t st);   // 2 kernels: select, pass
cudaError_t launch_resolve(const DeviceWorld& W, double dt, int do_euler, int prepare_blocks, int blocks, cudaStream_t st);
struct Mirror { uint32_t* ids; IcxTransform* T; IcxPhysics* P; IcxColliderComponent* C; uint8_t* hasP; };
void launch_import_kernel(const DeviceWorld& W, Mirror M, const uint32_t* index, const uint32_t* ids, const IcxTransform* T,
                          const IcxPhysics* P, const IcxColliderComponent* C, const uint8_t* hasP, const uint8_t* ts,
                          const uint16_t* tc, uint32_t count, int full, int have_C, uint32_t epoch, cudaStream_t st);
void launch_wire_messages(const DeviceWorld& W, Mirror M, const IcxWireLayout& lay, uint32_t coherency, int have_C, uint32_t* out, uint32_t n, cudaStream_t st);
void launch_motion_export(const DeviceWorld& W, IcxMotionState* out, uint32_t n, cudaStream_t st);
void launch_export_kernel(const DeviceWorld& W, Mirror M, int have_T, int have_P, int have_C, uint8_t* ts, uint16_t* tc,
                          uint32_t n, cudaStream_t st);

void launch_gather_rows(const DeviceWorld& W, Mirror M, const uint32_t* idx, uint32_t count, int mode, uint32_t* o_ids, IcxTransform* o_T,
                        IcxPhysics* o_P, IcxColliderComponent* o_C, uint8_t* o_hasP, uint8_t* o_ts, uint16_t* o_tc, cudaStream_t st);
void launch_extract_ts(const DeviceWorld& W, const uint32_t* index, const uint8_t* ts, uint32_t count, cudaStream_t st);

void launch_packet_structure(const IcxColliderComponent* C, uint32_t n, uint32_t* scan_buf, uint32_t* partial, uint8_t* mask, uint32_t* first,
                             cudaStream_t st);

struct DeviceBuffer {
    void* ptr = nullptr; size_t bytes = 0;
    cudaError_t ensure(size_t n) {
        if (n <= bytes) return cudaSuccess;
        if (ptr) cudaFree(ptr);
        ptr = nullptr; bytes = 0;
        cudaError_t e = cudaMalloc(&ptr, n);
        if (e == cudaSuccess) bytes = n;
        return e;
    }
    void release() { if (ptr) cudaFree(ptr); ptr = nullptr; bytes = 0; }
};

} // namespace icx

struct IcxWorld {
    IcxConfig cfg;
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    icx::DeviceWorld W;                 // device pointers + parameters (in/out already oriented)
    std::vector<icx::DeviceBuffer> pool;
    uint32_t cap_entities = 0, cap_packets = 0, cap_buckets = 0;
    uint32_t n_entities = 0, n_packets = 0;
    std::vector<uint8_t> active_mask;   // per entity: bit s = collider slot s active (packet structure); see icx_host_masks()
    bool masks_on_host = false;         // active_mask mirrors d_mask (fetched on demand: only update / clone / extract need it)
    icx::DeviceBuffer d_mask, st_scan;  // device copy of active_mask; scan scratch of the packet-structure pass
    std::vector<uint32_t> h_ids;        // host copy of the entity ids (clone / extract bookkeeping)
    icx::DeviceBuffer cl_idx, cl_ids, cl_T, cl_P, cl_C, cl_hasP, cl_ts, cl_tc;     // compact staging of clone / extract
    // staging (device) for AoS import/export, pinned host mirror for counters
    icx::DeviceBuffer mir_ids, mir_T, mir_P, mir_C, mir_hasP;       // device AoS mirror
    icx::DeviceBuffer st_ts, st_tc, st_index, st_T, st_P, st_C;      // staging for tstate and partial updates
    bool have_P = false, have_C = false;
    uint32_t c_epoch = 1;               // bumped whenever some ColliderComponent changes (upload / update / extract / replayed delta); DeviceWorld::c_ver
    // collider deltas a replay met while their entity was divergent: kept until the entity takes records again (icx_timeline.cu)
    icx::DeviceBuffer pend_rows, pend_flag, pend_list, pend_count;
    bool pend_any = false;
    // static triangle mesh (extension, icx_mesh.cu): device arrays + the pass's contact list
    icx::DeviceBuffer mesh_tri, mesh_cell_start, mesh_cell_tri, mesh_contacts, mesh_counter, mesh_active;
    int mesh_sms = 148;
    icx::DeviceMesh mesh = {};
    bool mesh_ran = false;              // a mesh pass ran in the last collision pass (its contacts are the ones to fetch)
    cudaEvent_t mesh_ev0 = nullptr, mesh_ev1 = nullptr;
    uint32_t* h_counters = nullptr;     // pinned
    uint32_t* h_flags = nullptr;        // pinned
    IcxStats stats;
    float last_ms = 0.0f;
    // optional per-phase device timing (icx_profile_enable): integrate_prep | grid sort | pairs | resolve
    bool profile = false;
    cudaEvent_t pev[ICX_N_PHASES + 1] = {};
    double phase_ms[ICX_N_PHASES] = {};
    uint64_t phase_ticks = 0;
    uint64_t totals[6] = {};            // since icx_profile_enable: ticks, candidate pairs, contacts, rounds, attempts, dirty bodies (summed over ticks)
    uint64_t launches = 0;              // kernels launched by icx_step* since creation
    int resolve_blocks = 0, prepare_blocks = 0, narrow_blocks = 0;
    std::string error;
    bool have_results = false;
    // behaviours on the device (icx_behaviors.cu): ProjectileComponent rows, their entity rows, condemned list + count
    icx::DeviceBuffer pj_proj, pj_index, pj_condemned, pj_count;
    uint32_t n_projectiles = 0;
    bool pj_updated = false;
    cudaEvent_t pj_ev0 = nullptr, pj_ev1 = nullptr;
};

// Grid buckets for the colliders actually active: the table is allocated for the capacity (2 colliders per entity), but the
// per-tick memsets and scans run over n_buckets, so a world with one collider per entity should not pay for twice that.
#ifndef ICX_BUCKET_FACTOR
#define ICX_BUCKET_FACTOR 2u
#endif
inline void icx_size_buckets(IcxWorld* w)
{
    uint32_t nb = 1024u;
    while (nb < ICX_BUCKET_FACTOR * w->n_packets && nb < w->cap_buckets) nb <<= 1;
    w->W.n_buckets = nb < w->cap_buckets ? nb : w->cap_buckets;
}

// make w->active_mask valid (one small device-to-host copy after an upload); ICX_OK or ICX_E_CUDA
int icx_host_masks(IcxWorld* w);
int icx_mesh_pass(IcxWorld* w);       // after a successful collision pass (state in W.in)
void icx_mesh_release(IcxWorld* w);
