// TEST TOOL: runs the product's box_box_intersect on host and device for one pair and prints both.
#include <cstdio>
#include "../../intercession_b200/csrc/icx_response.cuh"
using namespace icx;

__host__ __device__ static Shape mk(float ox, float oy, float oz, float sx, float sy, float sz)
{
    Q4 q; q.w = 1; q.x = q.y = q.z = 0;
    F4 P = model_transform(v3(ox, oy, oz), q, v3(sx, sy, sz));
    F4 L = model_transform(v3(0, 0, 0), q, v3(1, 1, 1));
    F4 f = mul(P, L);
    Shape s;
    s.M.c[0] = v3(f.c[0].x, f.c[0].y, f.c[0].z); s.M.c[1] = v3(f.c[1].x, f.c[1].y, f.c[1].z); s.M.c[2] = v3(f.c[2].x, f.c[2].y, f.c[2].z);
    s.M.t = v3(f.c[3].x, f.c[3].y, f.c[3].z);
    s.inv = inverse(mat3_of(s.M)); s.type = 1;
    return s;
}
__host__ __device__ static void run(const char* tag)
{
    Shape A = mk(-5, 0, 0, 2, 12, 12), B = mk(0, -5, 0, 12, 2, 12);
    Hit h; h.point = v3(0, 0, 0); h.normal = v3(0, 0, 0); h.depth = 0;
    V3 mA[8], mB[ICX_MAXV];
    bool hit = box_box_intersect(A, B, h);
    printf("%s hit=%d n=(%g %g %g) d=%g p=(%g %g %g)\n", tag, (int)hit, h.normal.x, h.normal.y, h.normal.z, h.depth, h.point.x, h.point.y, h.point.z);
    int nA = build_contact_manifold(A.M, -h.normal, mA), nB = build_contact_manifold(B.M, h.normal, mB);
    printf("%s nA=%d nB=%d\n", tag, nA, nB);
    for (int i = 0; i < nA; i++) printf("%s  A%d (%g %g %g)\n", tag, i, mA[i].x, mA[i].y, mA[i].z);
    for (int i = 0; i < nB; i++) printf("%s  B%d (%g %g %g)\n", tag, i, mB[i].x, mB[i].y, mB[i].z);
    pseudo_clip_polyhedra(mA, nA, mB, nB, h.normal);
    printf("%s clipped nB=%d\n", tag, nB);
    for (int i = 0; i < nB; i++) printf("%s  C%d (%g %g %g)\n", tag, i, mB[i].x, mB[i].y, mB[i].z);
}
__global__ void k() { run("DEV"); }
int main() { run("HOST"); k<<<1, 1>>>(); cudaError_t e = cudaDeviceSynchronize(); printf("cuda: %s\n", cudaGetErrorString(e)); return 0; }
