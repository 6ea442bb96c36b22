// dev: cost of getting pinned host memory three ways (cudaHostAlloc, malloc + cudaHostRegister, 2 MB-aligned + MADV_HUGEPAGE +
// cudaHostRegister), and of a pageable vs pinned H2D copy.   nvcc -O2 -o /tmp/pin_bench scripts/dev/pin_bench.cu
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cuda_runtime.h>
#include <sys/mman.h>
static double now() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
int main() {
    cudaFree(0);
    void* d; cudaMalloc(&d, 64 << 20);
    for (size_t mb : {8, 32, 64}) {
        const size_t n = mb << 20;
        double t = now(); void* p; cudaHostAlloc(&p, n, cudaHostAllocPortable); double a = now() - t;
        t = now(); cudaMemcpy(d, p, n, cudaMemcpyHostToDevice); double c1 = now() - t;
        t = now(); cudaFreeHost(p); double f = now() - t;
        t = now(); void* q = aligned_alloc(4096, n); memset(q, 0, n); double touch = now() - t;
        t = now(); cudaMemcpy(d, q, n, cudaMemcpyHostToDevice); double c2 = now() - t;
        t = now(); cudaError_t e = cudaHostRegister(q, n, cudaHostRegisterPortable); double r = now() - t;
        t = now(); cudaHostUnregister(q); double u = now() - t; free(q);
        t = now(); void* h = aligned_alloc(2 << 20, n); madvise(h, n, MADV_HUGEPAGE); memset(h, 0, n); double touch2 = now() - t;
        t = now(); cudaError_t e2 = cudaHostRegister(h, n, cudaHostRegisterPortable); double r2 = now() - t;
        cudaHostUnregister(h); free(h);
        printf("%zu MB: cudaHostAlloc %.2f ms (free %.2f), pinned H2D %.2f | malloc+touch %.2f, pageable H2D %.2f, register %.2f (%d) unregister %.2f | THP touch %.2f register %.2f (%d)\n",
               mb, a, f, c1, touch, c2, r, (int)e, u, touch2, r2, (int)e2);
    }
    FILE* fp = fopen("/sys/kernel/mm/transparent_hugepage/enabled", "r"); char buf[128] = {0}; if (fp) { fgets(buf, 127, fp); fclose(fp); } printf("THP: %s", buf);
    return 0;
}
