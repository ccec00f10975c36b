// C++ host, no Python, no torch: configs 4 and 5 across GPUs through include/icx_nccl.h with the caller's own ncclComm_t
// (one process, one communicator per device: ncclCommInitAll; with a single GPU the same calls run at world size 1).
//   config 4  every device holds W world-line candidates x 1024 bodies, ticks them, scans against a base record,
//             ONE icx_scores_allgather returns every rank's verdicts to every rank
//   config 5  device 0 records a tick and sends the slot to the last device (ncclSend / ncclRecv inside a group), which imports
//             and replays it: the receiver's state equals the sender's, bit for bit
// Prints one line per check; tests/test_host_cpp.py asserts on them.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <vector>
#include <dlfcn.h>
#include <pthread.h>
#include <cuda_runtime_api.h>
#include "../../include/icx_nccl.h"

typedef int (*fn_init_all)(ncclComm_t*, int, const int*);
typedef int (*fn_destroy)(ncclComm_t);

#define CK(call) do { int rc_ = (call); if (rc_ != ICX_OK) { std::printf("FAIL %s -> %d\n", #call, rc_); return 1; } } while (0)

static void make_scene(uint32_t n, uint32_t seed, std::vector<IcxEntity>& ids, std::vector<IcxTransform>& T, std::vector<IcxPhysics>& P,
                       std::vector<IcxColliderComponent>& C, uint32_t link)
{
    ids.resize(n); T.assign(n, IcxTransform{}); P.assign(n, IcxPhysics{}); C.assign(n, IcxColliderComponent{});
    uint32_t s = seed * 2654435761u + 12345u;
    auto rnd = [&]() { s = s * 1664525u + 1013904223u; return (float)(s >> 8) / 16777216.0f; };
    const uint32_t m = (uint32_t)std::ceil(std::cbrt((double)n));
    for (uint32_t i = 0; i < n; i++) {
        ids[i] = (14u << 28) | (i << 4) | link;
        T[i].origin[0] = 1.3f * (float)(i % m) + 0.2f * rnd(); T[i].origin[1] = 1.3f * (float)((i / m) % m) + 0.2f * rnd(); T[i].origin[2] = 1.3f * (float)(i / (m * m)) + 0.2f * rnd();
        T[i].orientation[0] = 1.0f; T[i].scale[0] = T[i].scale[1] = T[i].scale[2] = 0.6f + 0.4f * rnd();
        P[i].linearDrag = 0.01f; P[i].angularDrag = 0.01f; P[i].collisionAngularDrag = 0.03f; P[i].mass = 5.0f; P[i].lockedOrientation[0] = 1.0f;
        P[i].velocity[0] = 2.0f * rnd() - 1.0f; P[i].velocity[1] = 2.0f * rnd() - 1.0f; P[i].velocity[2] = 2.0f * rnd() - 1.0f;
        for (int k = 0; k < 2; k++) {
            IcxCollider& c = C[i].colliders[k];
            c.colliderType = (i & 1) ? ICX_COLLIDER_BOX : ICX_COLLIDER_SPHERE; c.collisionType = ICX_COLLISION_RIGID; c.isActive = k == 0;
            c.localTransform.orientation[0] = 1.0f; c.localTransform.scale[0] = c.localTransform.scale[1] = c.localTransform.scale[2] = 1.0f;
            c.inheritOrientation = 1; c.influenceOrientation = 1; c.minParametricValue = 1.0f; c.staticFriction = 0.45f; c.dynamicFriction = 0.45f;
            c.restitution = 0.7f; c.stiffness = 100.0f; c.damping = 10.0f; c.restLength = 0.1f;
        }
    }
}

int main()
{
    if (icx_device_count() == 0) { std::printf("no CUDA device: no CPU fallback\n"); return 2; }
    if (!icx_nccl_available()) { std::printf("no NCCL library in this process\n"); return 3; }
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    fn_init_all init_all = h ? (fn_init_all)dlsym(h, "ncclCommInitAll") : nullptr;
    fn_destroy destroy = h ? (fn_destroy)dlsym(h, "ncclCommDestroy") : nullptr;
    if (!init_all || !destroy) { std::printf("ncclCommInitAll not found\n"); return 3; }
    const int R = icx_device_count() >= 2 ? 2 : 1;
    int devs[2] = {0, 1};
    ncclComm_t comm[2] = {nullptr, nullptr};
    if (init_all(comm, R, devs) != 0) { std::printf("FAIL ncclCommInitAll\n"); return 1; }
    std::printf("ranks %d\n", R);

    // ---------------- config 4: sharded world-line candidates + ONE all-gather ----------------
    const uint32_t per = 1024, wl_per_rank = 3;
    std::vector<IcxEntity> ids; std::vector<IcxTransform> T; std::vector<IcxPhysics> P; std::vector<IcxColliderComponent> C;
    IcxWorld* base[2]; IcxWorld* cand[2]; IcxTimeline* tlb[2];
    for (int r = 0; r < R; r++) {
        IcxConfig cfg; icx_config_default(&cfg); cfg.device = devs[r]; cfg.max_entities = per; cfg.host_timeslice = ICX_NULL_TIMESLICE;
        CK(icx_world_create(&cfg, &base[r]));
        make_scene(per, 7, ids, T, P, C, 1);
        CK(icx_bodies_upload(base[r], per, ids.data(), T.data(), P.data(), nullptr, C.data(), nullptr, nullptr));
        CK(icx_timeline_create(devs[r], 8, per, &tlb[r]));
        cfg.max_entities = per * wl_per_rank; cfg.worldline_stride = per;
        CK(icx_world_create(&cfg, &cand[r]));
        std::vector<IcxEntity> ci; std::vector<IcxTransform> cT; std::vector<IcxPhysics> cP; std::vector<IcxColliderComponent> cC;
        for (uint32_t w = 0; w < wl_per_rank; w++) {
            make_scene(per, 7, ids, T, P, C, 1);
            const uint32_t global_wl = (uint32_t)r * wl_per_rank + w;
            if (global_wl % 2 == 1) { P[global_wl].velocity[0] += 40.0f; }          // odd candidates diverge: one body flies off
            for (uint32_t i = 0; i < per; i++) ids[i] = (14u << 28) | ((w * per + i) << 4) | 1u;
            ci.insert(ci.end(), ids.begin(), ids.end()); cT.insert(cT.end(), T.begin(), T.end()); cP.insert(cP.end(), P.begin(), P.end()); cC.insert(cC.end(), C.begin(), C.end());
        }
        CK(icx_bodies_upload(cand[r], per * wl_per_rank, ci.data(), cT.data(), cP.data(), nullptr, cC.data(), nullptr, nullptr));
    }
    for (int t = 0; t < 6; t++)
        for (int r = 0; r < R; r++) {
            CK(icx_step(base[r], 1.0 / 45)); CK(icx_record_ex(base[r], tlb[r], (uint16_t)(t + 1), ICX_RECORD_KEEP_LINK));
            CK(icx_step(cand[r], 1.0 / 45)); CK(icx_divergence_scan_enqueue(cand[r], tlb[r], (uint16_t)(t + 1), wl_per_rank));
        }
    std::vector<uint8_t> ever[2]; std::vector<uint32_t> cnt[2]; std::vector<double> score[2];
    for (int r = 0; r < R; r++) { ever[r].assign(wl_per_rank * R, 9); cnt[r].assign(wl_per_rank * R, 0); score[r].assign(wl_per_rank * R, 0.0); }
    if (R == 1) CK(icx_scores_allgather(tlb[0], wl_per_rank, wl_per_rank, 1, comm[0], ever[0].data(), cnt[0].data(), score[0].data()));
    else {
        // two devices, one thread: the second rank's call would block the first -> run them from two host threads
        struct Arg { IcxTimeline* tl; ncclComm_t c; uint8_t* e; uint32_t* n; double* s; int rc; } a[2];
        for (int r = 0; r < 2; r++) a[r] = {tlb[r], comm[r], ever[r].data(), cnt[r].data(), score[r].data(), -1};
        pthread_t th[2];
        auto run = [](void* p) -> void* { Arg* x = (Arg*)p; x->rc = icx_scores_allgather(x->tl, 3, 3, 2, x->c, x->e, x->n, x->s); return nullptr; };
        for (int r = 0; r < 2; r++) pthread_create(&th[r], nullptr, run, &a[r]);
        for (int r = 0; r < 2; r++) pthread_join(th[r], nullptr);
        if (a[0].rc != ICX_OK || a[1].rc != ICX_OK) { std::printf("FAIL allgather %d %d\n", a[0].rc, a[1].rc); return 1; }
    }
    int ok4 = 1;
    for (int r = 0; r < R; r++)
        for (uint32_t k = 0; k < wl_per_rank * R; k++) {
            const int want = (k % 2 == 1) ? 1 : 0;
            if (ever[r][k] != want) ok4 = 0;
            if (want && !(cnt[r][k] >= 1 && score[r][k] > 2.0)) ok4 = 0;
            if (r > 0 && (ever[r][k] != ever[0][k] || cnt[r][k] != cnt[0][k] || score[r][k] != score[0][k])) ok4 = 0;
        }
    std::printf("config4 allgather %s worldlines %u divergent %d\n", ok4 ? "ok" : "MISMATCH", wl_per_rank * R, (int)(wl_per_rank * R / 2));

    // ---------------- config 5: one snapshot future -> past over ncclSend / ncclRecv ----------------
    const int src = 0, dst = R - 1;
    const uint32_t n = 2000;
    IcxWorld* wf; IcxWorld* wp; IcxTimeline* tf; IcxTimeline* tp;
    IcxConfig cfg; icx_config_default(&cfg); cfg.max_entities = n;
    cfg.device = devs[src]; cfg.host_timeslice = 0; CK(icx_world_create(&cfg, &wf));
    cfg.device = devs[dst]; cfg.host_timeslice = 1; CK(icx_world_create(&cfg, &wp));
    make_scene(n, 3, ids, T, P, C, 0);
    for (uint32_t i = 0; i < n; i++) ids[i] = (i << 4) | 0u;
    CK(icx_bodies_upload(wf, n, ids.data(), T.data(), P.data(), nullptr, C.data(), nullptr, nullptr));
    std::vector<IcxEntity> pid(n); for (uint32_t i = 0; i < n; i++) pid[i] = ids[i] + 1;       // the past knows the entity at chain-link + 1
    std::vector<IcxTransform> T0 = T; for (auto& t : T0) t.origin[1] += 5.0f;                 // the past starts somewhere else
    CK(icx_bodies_upload(wp, n, pid.data(), T0.data(), P.data(), nullptr, C.data(), nullptr, nullptr));
    CK(icx_timeline_create(devs[src], 4, n, &tf)); CK(icx_timeline_create(devs[dst], 4, n, &tp));
    CK(icx_step_n(wf, 1.0 / 45, 5));
    CK(icx_record(wf, tf, 105));
    const size_t bytes = ICX_SNAPSHOT_CAPACITY(n, n);
    void* bs; void* br; cudaSetDevice(devs[src]); cudaMalloc(&bs, bytes); cudaSetDevice(devs[dst]); cudaMalloc(&br, bytes);
    // one stream per communicator inside a group (with a single rank both calls use the same communicator); the replay on the
    // receiving world's own stream is ordered behind the import by the slot's event
    void* recv_stream = R == 1 ? icx_world_stream(wf) : icx_world_stream(wp);
    CK(icx_nccl_group_start());
    CK(icx_timeline_slot_send(tf, 105, comm[src], dst, bs, bytes, icx_world_stream(wf)));
    CK(icx_timeline_slot_recv(tp, 105, n, comm[dst], src, br, bytes, recv_stream));
    CK(icx_nccl_group_end());
    int applied = 0;
    CK(icx_replay(wp, tp, 105, ICX_REPLAY_SERVER, &applied));
    std::vector<IcxTransform> Ta(n), Tb(n); std::vector<IcxPhysics> Pa(n), Pb(n);
    CK(icx_bodies_download(wf, n, Ta.data(), Pa.data(), nullptr, nullptr, nullptr));
    CK(icx_bodies_download(wp, n, Tb.data(), Pb.data(), nullptr, nullptr, nullptr));
    uint64_t mism = 0; CK(icx_timeline_id_mismatches(tp, &mism));
    const int same = std::memcmp(Ta.data(), Tb.data(), n * sizeof(IcxTransform)) == 0 && std::memcmp(Pa.data(), Pb.data(), n * sizeof(IcxPhysics)) == 0;
    std::printf("config5 send/recv %s applied %d id_mismatches %llu bytes %zu\n", same ? "ok" : "MISMATCH", applied, (unsigned long long)mism, bytes);
    for (int r = 0; r < R; r++) destroy(comm[r]);
    return (ok4 && same && applied == 1 && mism == 0) ? 0 : 1;
}
