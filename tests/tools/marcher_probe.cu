// Debug tool: evaluates the device Mandelbulb routines along one ray on the GPU and prints the
// bits, to be diffed against the same source compiled for the host.
#include <cstdio>
#include <cstring>
#include "../../raysnail_b200/csrc/device/rt_scene.cuh"
using namespace rsd;
__global__ void probe(Ray ray, double* r_out, double* dr_out, int* in_out, double* t_out) {
    const double len = length(ray.d);
    const V3 direction = ray.d * (1.0 / len);
    const V3 df = direction * 0.05;
    V3 v = ray.o;
    t_out[2] = len; t_out[3] = direction.x; t_out[4] = direction.y; t_out[5] = direction.z; t_out[6] = df.x; t_out[7] = df.y; t_out[8] = df.z;
    for (int i = 0; i < 200; ++i) {
        if (i == 170) { t_out[9] = v.x; t_out[10] = v.y; t_out[11] = v.z; }
        double r, dr;
        in_out[i] = marcher_iterate(v, 100, r, dr) ? 1 : 0;
        r_out[i] = r; dr_out[i] = dr;
        v = v + df;
    }
    rs_node n;
    memset(&n, 0, sizeof(n));
    n.p[0] = 100.0;
    HitRec h;
    h.t1 = 0.0;
    bool ok = hit_marcher<false>(n, ray, 1e-4, 1e300, h);
    t_out[0] = ok ? h.t1 : -1.0;
    ok = hit_marcher<true>(n, ray, 1e-4, 1e300, h);
    t_out[1] = ok ? h.t1 : -1.0;
}
int main() {
    const char* hex[6] = {"0x1.23333333333334p+3", "-0x1.30a3d70a3d70ap+0", "0x1.0cccccccccccdp+1", "0", "0", "0"};
    (void)hex;
    double v[6];
    for (int i = 0; i < 6; ++i) if (scanf("%la", &v[i]) != 1) return 1;
    Ray ray; ray.o.x = v[0]; ray.o.y = v[1]; ray.o.z = v[2]; ray.d.x = v[3]; ray.d.y = v[4]; ray.d.z = v[5]; ray.time = 0.0;
    double *r, *dr, *t; int* in;
    cudaMallocManaged(&r, 200 * 8); cudaMallocManaged(&dr, 200 * 8); cudaMallocManaged(&in, 200 * 4); cudaMallocManaged(&t, 16 * 8);
    probe<<<1, 1>>>(ray, r, dr, in, t);
    if (cudaDeviceSynchronize() != cudaSuccess) { printf("cuda error\n"); return 2; }
    for (int i = 0; i < 200; ++i) printf("%d %d %a %a\n", i, in[i], r[i], dr[i]);
    printf("t %a %a (%.17g %.17g)\n", t[0], t[1], t[0], t[1]);
    for (int i = 2; i < 12; ++i) printf("aux %d %a\n", i, t[i]);
    return 0;
}
