// C ABI, part 5: the static triangle mesh (extension without a counterpart in the reference, see include/icx.h and
// icx_mesh.cuh).  Host side: the triangle soup as 3 float4 per triangle, a uniform grid over it (CSR, ascending triangle
// indices per cell) built once at upload, the pass after every collision pass, and the ordered contact fetch.
#include <algorithm>
#include <cmath>
#include <cstring>
#include "icx_host.h"

using namespace icx;

namespace {
int mfail(IcxWorld* w, int code, const std::string& msg) { if (w) w->error = msg; return code; }
#define CKM(w, call) do { cudaError_t _e = (call); if (_e != cudaSuccess) { (w)->error = std::string(#call) + ": " + cudaGetErrorString(_e); return ICX_E_CUDA; } } while (0)
}

void icx_mesh_release(IcxWorld* w)
{
    w->mesh_tri.release(); w->mesh_cell_start.release(); w->mesh_cell_tri.release(); w->mesh_contacts.release(); w->mesh_counter.release(); w->mesh_active.release();
    if (w->mesh_ev0) cudaEventDestroy(w->mesh_ev0);
    if (w->mesh_ev1) cudaEventDestroy(w->mesh_ev1);
    w->mesh_ev0 = w->mesh_ev1 = nullptr;
    w->mesh = DeviceMesh{}; w->mesh_ran = false;
}

int icx_mesh_pass(IcxWorld* w)
{
    w->mesh_ran = false;
    if (!w->mesh.n_tri || !w->n_entities) return ICX_OK;
    cudaStream_t st = w->stream;
    CKM(w, cudaMemsetAsync(w->mesh_counter.ptr, 0, 16, st));
    CKM(w, cudaEventRecord(w->mesh_ev0, st));
    launch_mesh_pass(w->W, w->mesh, w->mesh_sms, st);
    CKM(w, cudaEventRecord(w->mesh_ev1, st));
    CKM(w, cudaGetLastError());
    w->launches += 2;
    w->mesh_ran = true;     // the contact list holds one contact per entity of capacity to start with; an overflow is reported by the fetch, which enlarges it
    return ICX_OK;
}

extern "C" {

int icx_mesh_upload(IcxWorld* w, const float* vertices, uint32_t n_vertices, const uint32_t* triangles, uint32_t n_triangles, const IcxMeshMaterial* material)
{
    if (!w) return ICX_E_INVALID;
    cudaSetDevice(w->device);
    CKM(w, cudaStreamSynchronize(w->stream));
    if (n_triangles == 0) { icx_mesh_release(w); return ICX_OK; }
    if (!vertices || !triangles || n_vertices == 0) return mfail(w, ICX_E_INVALID, "mesh_upload: null vertices / triangles");
    std::vector<float4> tri((size_t)n_triangles * 3);
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    double extent = 0.0;
    for (uint32_t t = 0; t < n_triangles; t++) {
        float tlo[3] = {INFINITY, INFINITY, INFINITY}, thi[3] = {-INFINITY, -INFINITY, -INFINITY};
        for (int k = 0; k < 3; k++) {
            const uint32_t v = triangles[3 * t + k];
            if (v >= n_vertices) return mfail(w, ICX_E_INVALID, "mesh_upload: vertex index out of range");
            const float* p = vertices + 3 * (size_t)v;
            if (!(std::isfinite(p[0]) && std::isfinite(p[1]) && std::isfinite(p[2]))) return mfail(w, ICX_E_INVALID, "mesh_upload: non-finite vertex");
            tri[3 * (size_t)t + k] = make_float4(p[0], p[1], p[2], 0.0f);
            for (int a = 0; a < 3; a++) { tlo[a] = std::min(tlo[a], p[a]); thi[a] = std::max(thi[a], p[a]); }
        }
        for (int a = 0; a < 3; a++) { lo[a] = std::min(lo[a], tlo[a]); hi[a] = std::max(hi[a], thi[a]); }
        extent += std::max(std::max(thi[0] - tlo[0], thi[1] - tlo[1]), thi[2] - tlo[2]);
    }
    // cell = twice the mean triangle extent, at most 128 cells per axis
    float cell = (float)(2.0 * extent / n_triangles);
    const float span = std::max(std::max(hi[0] - lo[0], hi[1] - lo[1]), hi[2] - lo[2]);
    if (!(cell > 0.0f)) cell = span > 0.0f ? span : 1.0f;
    cell = std::max(cell, span / 128.0f);
    DeviceMesh M = {};
    const float pad = 1.0e-3f * cell;                       // registration is conservative; the pass tests the exact bounds
    for (int a = 0; a < 3; a++) { M.lo[a] = lo[a] - pad; M.dim[a] = std::max(1, std::min(128, (int)std::floor((hi[a] + pad - M.lo[a]) / cell) + 1)); }
    M.inv_cell = 1.0f / cell;
    const size_t n_cells = (size_t)M.dim[0] * M.dim[1] * M.dim[2];
    std::vector<uint32_t> start(n_cells + 1, 0);
    auto cells_of = [&](uint32_t t, int* c0, int* c1) {
        for (int a = 0; a < 3; a++) {
            const float4 A = tri[3 * (size_t)t], B = tri[3 * (size_t)t + 1], C = tri[3 * (size_t)t + 2];
            const float pa[3] = {A.x, A.y, A.z}, pb[3] = {B.x, B.y, B.z}, pc[3] = {C.x, C.y, C.z};
            const float tl = std::min(std::min(pa[a], pb[a]), pc[a]) - pad, th = std::max(std::max(pa[a], pb[a]), pc[a]) + pad;
            c0[a] = std::max(0, std::min(M.dim[a] - 1, (int)std::floor((tl - M.lo[a]) * M.inv_cell)));
            c1[a] = std::max(0, std::min(M.dim[a] - 1, (int)std::floor((th - M.lo[a]) * M.inv_cell)));
        }
    };
    for (int fill = 0; fill < 2; fill++) {
        std::vector<uint32_t> at;
        std::vector<uint32_t> list;
        if (fill) { at.assign(start.begin(), start.end() - 1); list.resize(start[n_cells]); }
        for (uint32_t t = 0; t < n_triangles; t++) {        // ascending t: every cell's list comes out ascending
            int c0[3], c1[3]; cells_of(t, c0, c1);
            for (int z = c0[2]; z <= c1[2]; z++) for (int y = c0[1]; y <= c1[1]; y++) for (int x = c0[0]; x <= c1[0]; x++) {
                const size_t cell_i = ((size_t)z * M.dim[1] + y) * M.dim[0] + x;
                if (fill) list[at[cell_i]++] = t; else start[cell_i + 1]++;
            }
        }
        if (!fill) { for (size_t c = 0; c < n_cells; c++) start[c + 1] += start[c]; continue; }
        CKM(w, w->mesh_tri.ensure(tri.size() * sizeof(float4)));
        CKM(w, w->mesh_cell_start.ensure(start.size() * 4));
        CKM(w, w->mesh_cell_tri.ensure(std::max<size_t>(list.size(), 1) * 4));
        CKM(w, cudaMemcpy(w->mesh_tri.ptr, tri.data(), tri.size() * sizeof(float4), cudaMemcpyHostToDevice));
        CKM(w, cudaMemcpy(w->mesh_cell_start.ptr, start.data(), start.size() * 4, cudaMemcpyHostToDevice));
        if (!list.empty()) CKM(w, cudaMemcpy(w->mesh_cell_tri.ptr, list.data(), list.size() * 4, cudaMemcpyHostToDevice));
    }
    M.max_contacts = std::max<uint32_t>(4096u, w->cap_entities);           // grows when a pass overflows it (icx_mesh_contacts_fetch)
    CKM(w, w->mesh_contacts.ensure((size_t)M.max_contacts * sizeof(IcxMeshContact)));
    CKM(w, w->mesh_counter.ensure(16));
    CKM(w, cudaMemset(w->mesh_counter.ptr, 0, 16));
    CKM(w, w->mesh_active.ensure((size_t)std::max<uint32_t>(w->cap_entities, 1u) * 4));
    cudaDeviceGetAttribute(&w->mesh_sms, cudaDevAttrMultiProcessorCount, w->device);
    if (!w->mesh_ev0) { CKM(w, cudaEventCreate(&w->mesh_ev0)); CKM(w, cudaEventCreate(&w->mesh_ev1)); }
    M.tri = (const float4*)w->mesh_tri.ptr; M.cell_start = (const uint32_t*)w->mesh_cell_start.ptr; M.cell_tri = (const uint32_t*)w->mesh_cell_tri.ptr;
    M.n_tri = n_triangles;
    M.staticFriction = material ? material->staticFriction : 0.45f; M.dynamicFriction = material ? material->dynamicFriction : 0.45f;
    M.restitution = material ? material->restitution : 0.70f;
    M.contacts = (IcxMeshContact*)w->mesh_contacts.ptr; M.counter = (uint32_t*)w->mesh_counter.ptr; M.active = (uint32_t*)w->mesh_active.ptr;
    w->mesh = M; w->mesh_ran = false;
    return ICX_OK;
}

int icx_mesh_contacts_fetch(IcxWorld* w, IcxMeshContact* out, size_t cap, size_t* n_out)
{
    if (!w) return ICX_E_INVALID;
    if (n_out) *n_out = 0;
    if (!w->mesh_ran) return ICX_OK;
    cudaSetDevice(w->device);
    uint32_t cnt[2] = {0, 0};
    CKM(w, cudaMemcpyAsync(cnt, w->mesh_counter.ptr, 8, cudaMemcpyDeviceToHost, w->stream));
    CKM(w, cudaStreamSynchronize(w->stream));
    if (cnt[1] || cnt[0] > w->mesh.max_contacts) {
        // the pass itself is complete (states are right), its contact list is not: make room for the next pass and say so
        const uint32_t want = std::max(cnt[0] + cnt[0] / 2u, w->mesh.max_contacts * 2u);
        w->mesh_contacts.release();
        CKM(w, w->mesh_contacts.ensure((size_t)want * sizeof(IcxMeshContact)));
        w->mesh.contacts = (IcxMeshContact*)w->mesh_contacts.ptr; w->mesh.max_contacts = want;
        return mfail(w, ICX_E_CAPACITY, "mesh contact list overflowed in this pass (the states are complete); it has been enlarged for the next one");
    }
    if (n_out) *n_out = cnt[0];
    if (!out || !cnt[0]) return ICX_OK;
    std::vector<IcxMeshContact> all(cnt[0]);
    CKM(w, cudaMemcpy(all.data(), w->mesh_contacts.ptr, (size_t)cnt[0] * sizeof(IcxMeshContact), cudaMemcpyDeviceToHost));
    std::sort(all.begin(), all.end(), [](const IcxMeshContact& a, const IcxMeshContact& b) {      // the pass's own order: row, slot, triangle
        if (a.row != b.row) return a.row < b.row;
        if (a.slot != b.slot) return a.slot < b.slot;
        return a.triangle < b.triangle;
    });
    std::memcpy(out, all.data(), std::min<size_t>(cap, all.size()) * sizeof(IcxMeshContact));
    return ICX_OK;
}

int icx_mesh_kernel_us(IcxWorld* w, float* us)
{
    if (!w || !us) return ICX_E_INVALID;
    *us = 0.0f;
    if (!w->mesh_ran) return ICX_OK;
    cudaSetDevice(w->device);
    CKM(w, cudaEventSynchronize(w->mesh_ev1));
    float ms = 0.0f;
    CKM(w, cudaEventElapsedTime(&ms, w->mesh_ev0, w->mesh_ev1));
    *us = ms * 1e3f;
    return ICX_OK;
}

} // extern "C"
