// C ABI, part 3: behaviours on the device.
// ProjectileBehaviors::on_fixed_update (source/behaviors/projectile_behaviors.h:17-54): the reference
// walks its BehaviorsPackets on the host every fixed step and, per projectile, advances the lifetime,
// writes a pulsing TransformComponent::scale and condemns the entity once its lifetime is over.
// Here the ProjectileComponent rows live in HBM next to the world; one kernel does the whole pass and
// writes the scale into prm0, where the next tick reads it and from where k_export patches the AoS
// mirror for icx_bodies_download.
// HBM-bound: 40 B read + 8 B written per component row, 4 B index, 16 B read + 16 B written of prm0
// = 84 B per projectile.
#include <algorithm>
#include <string>
#include "icx_host.h"

using namespace icx;

namespace {

int bfail(IcxWorld* w, int code, const std::string& msg) { if (w) w->error = msg; return code; }
int bfail_cuda(IcxWorld* w, cudaError_t e, const char* what)
{
    if (w) w->error = std::string(what) + ": " + cudaGetErrorString(e);
    return ICX_E_CUDA;
}
#define CKB(w, call) do { cudaError_t _e = (call); if (_e != cudaSuccess) return bfail_cuda(w, _e, #call); } while (0)

#define PJ_TPB 256
#define PJ_ROW_WORDS 5                                   // doubles per ProjectileComponent

// One thread per projectile.  The block's 256 rows (10 240 contiguous bytes) are staged through shared
// memory with 128-bit loads, so the AoS rows cost full-width coalesced transactions; every thread then
// reads its five doubles from shared memory (stride 5 words of 8 B: conflict-free, 5 is odd).
// Arithmetic exactly as the reference evaluates it in double (no contraction: the library is built
// with -fmad=false), one rounding to float at the end.
__global__ void __launch_bounds__(PJ_TPB) k_projectile_update(double* __restrict__ proj, const uint32_t* __restrict__ index, uint32_t n, double dt,
                                                              float4* __restrict__ prm0,
                                                              uint32_t* __restrict__ condemned, uint32_t* __restrict__ n_condemned)
{
    __shared__ __align__(16) double rows[PJ_TPB * PJ_ROW_WORDS];
    const uint32_t tid = threadIdx.x, base = blockIdx.x * PJ_TPB;
    const uint32_t m = min((uint32_t)PJ_TPB, n - base);                           // rows of this block
    {
        const uint32_t n_vec = (m * PJ_ROW_WORDS + 1u) / 2u;                      // 16-byte words (the buffer is padded)
        const double2* src = reinterpret_cast<const double2*>(proj + (size_t)base * PJ_ROW_WORDS);
        double2* dst = reinterpret_cast<double2*>(rows);
        for (uint32_t v = tid; v < n_vec; v += PJ_TPB) dst[v] = src[v];
    }
    __syncthreads();
    bool dead = false;
    const uint32_t i = base + tid;
    if (tid < m) {
        const double* r = rows + tid * PJ_ROW_WORDS;
        const double life = r[0] + dt;                                            // :33
        const double maxLife = r[1], hz = r[2], scaleMax = r[3], scaleMin = r[4];
        const double s = sin(life * hz);                                          // :35
        const double amp = (scaleMax - scaleMin) / 2.0;                           // :38
        const float sc = (float)(s * amp + (scaleMin + 1.0));                     // :39
        proj[(size_t)i * PJ_ROW_WORDS] = life;
        const uint32_t e = index[i];
        float4 p0 = prm0[e];
        p0.x = sc; p0.y = sc; p0.z = sc;
        prm0[e] = p0;
        dead = life > maxLife;                                                    // :41
    }
    // condemned projectile numbers: one atomic per warp, order restored by the host (the list is short)
    const uint32_t mask = __ballot_sync(0xFFFFFFFFu, dead), lane = tid & 31u;
    if (mask) {
        uint32_t pos = 0;
        if (lane == (uint32_t)(__ffs(mask) - 1)) pos = atomicAdd(n_condemned, (uint32_t)__popc(mask));
        pos = __shfl_sync(0xFFFFFFFFu, pos, __ffs(mask) - 1);
        if (dead) condemned[pos + __popc(mask & ((1u << lane) - 1u))] = i;
    }
}

} // namespace

extern "C" {

int icx_projectiles_upload(IcxWorld* w, uint32_t n, const uint32_t* index, const IcxProjectile* proj)
{
    if (!w) return ICX_E_INVALID;
    if (n && (!index || !proj)) return bfail(w, ICX_E_INVALID, "projectiles: null index / components");
    if (n > w->n_entities) return bfail(w, ICX_E_INVALID, "projectiles: more projectiles than uploaded entities");
    for (uint32_t i = 0; i < n; i++) if (index[i] >= w->n_entities) return bfail(w, ICX_E_INVALID, "projectiles: index out of range");
    cudaSetDevice(w->device);
    w->n_projectiles = 0; w->pj_updated = false;
    if (!n) return ICX_OK;
    CKB(w, w->pj_proj.ensure((size_t)n * sizeof(IcxProjectile) + 16));
    CKB(w, w->pj_index.ensure((size_t)n * 4));
    CKB(w, w->pj_condemned.ensure((size_t)n * 4));
    CKB(w, w->pj_count.ensure(16));
    if (!w->pj_ev0) { CKB(w, cudaEventCreate(&w->pj_ev0)); CKB(w, cudaEventCreate(&w->pj_ev1)); }
    CKB(w, cudaMemcpyAsync(w->pj_proj.ptr, proj, (size_t)n * sizeof(IcxProjectile), cudaMemcpyHostToDevice, w->stream));
    CKB(w, cudaMemcpyAsync(w->pj_index.ptr, index, (size_t)n * 4, cudaMemcpyHostToDevice, w->stream));
    CKB(w, cudaStreamSynchronize(w->stream));
    w->n_projectiles = n;
    return ICX_OK;
}

int icx_projectiles_fixed_update(IcxWorld* w, double dt)
{
    if (!w) return ICX_E_INVALID;
    if (!(dt >= 0.0)) return bfail(w, ICX_E_INVALID, "projectiles: deltaTime < 0");          // assert, projectile_behaviors.h:24
    if (!w->n_projectiles) { w->pj_updated = false; return ICX_OK; }
    cudaSetDevice(w->device);
    cudaStream_t st = w->stream;
    CKB(w, cudaMemsetAsync(w->pj_count.ptr, 0, 4, st));
    CKB(w, cudaEventRecord(w->pj_ev0, st));
    k_projectile_update<<<(w->n_projectiles + PJ_TPB - 1) / PJ_TPB, PJ_TPB, 0, st>>>(
        (double*)w->pj_proj.ptr, (const uint32_t*)w->pj_index.ptr, w->n_projectiles, dt, w->W.prm0,
        (uint32_t*)w->pj_condemned.ptr, (uint32_t*)w->pj_count.ptr);
    CKB(w, cudaEventRecord(w->pj_ev1, st));
    CKB(w, cudaGetLastError());
    w->launches += 1;
    w->pj_updated = true;
    return ICX_OK;
}

int icx_projectiles_download(IcxWorld* w, uint32_t n, IcxProjectile* out)
{
    if (!w || (n && !out)) return ICX_E_INVALID;
    if (n > w->n_projectiles) return bfail(w, ICX_E_INVALID, "projectiles: n > uploaded projectiles");
    if (!n) return ICX_OK;
    cudaSetDevice(w->device);
    CKB(w, cudaMemcpyAsync(out, w->pj_proj.ptr, (size_t)n * sizeof(IcxProjectile), cudaMemcpyDeviceToHost, w->stream));
    CKB(w, cudaStreamSynchronize(w->stream));
    return ICX_OK;
}

int icx_projectiles_condemned(IcxWorld* w, uint32_t* out, size_t cap, size_t* n)
{
    if (!w || !n) return ICX_E_INVALID;
    *n = 0;
    if (!w->n_projectiles || !w->pj_updated) return ICX_OK;
    cudaSetDevice(w->device);
    uint32_t count = 0;
    CKB(w, cudaMemcpyAsync(&count, w->pj_count.ptr, 4, cudaMemcpyDeviceToHost, w->stream));
    CKB(w, cudaStreamSynchronize(w->stream));
    *n = count;
    if (!out || !count) return ICX_OK;
    if (cap < count) return bfail(w, ICX_E_CAPACITY, "projectiles: condemned list larger than the caller's buffer");
    CKB(w, cudaMemcpyAsync(out, w->pj_condemned.ptr, (size_t)count * 4, cudaMemcpyDeviceToHost, w->stream));
    CKB(w, cudaStreamSynchronize(w->stream));
    std::sort(out, out + count);
    return ICX_OK;
}

int icx_projectiles_kernel_ms(IcxWorld* w, float* ms)
{
    if (!w || !ms) return ICX_E_INVALID;
    if (!w->pj_updated) return bfail(w, ICX_E_STATE, "projectiles: no update has run");
    cudaSetDevice(w->device);
    CKB(w, cudaEventSynchronize(w->pj_ev1));
    CKB(w, cudaEventElapsedTime(ms, w->pj_ev0, w->pj_ev1));
    return ICX_OK;
}

} // extern "C"
