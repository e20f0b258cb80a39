// Probe: can launch n+1 of a one-CTA-per-SM kernel start on the SMs that launch n's first CTAs have left,
// while launch n's last CTAs are still running?  (The chained launch pays one pipeline fill per step because a step is
// a kernel)  Every CTA of launch L waits for the same CTA of launch L-1 through a
// release/acquire word in global memory, then works for base + idx * stagger microseconds.
//
// Variants: stream order only; programmatic dependent launch (PDL); PDL + cooperative attribute; PDL with an event record,
// an event wait, a small H2D copy or a D2D copy between the launches; two alternating streams.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o pdl_probe pdl_probe.cu && ./pdl_probe
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <algorithm>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("  %s -> %s\n", #x, cudaGetErrorString(e_)); return false; } } while (0)

__device__ __forceinline__ unsigned long long gtime()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

__global__ void __launch_bounds__(320) probe_kernel(unsigned long long *times, unsigned *flags, int launch, int nlaunch,
                                                    int trigger, unsigned base_ns, unsigned step_ns)
{
    extern __shared__ float smem[];
    if (trigger) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    const unsigned long long t0 = gtime();
    if (threadIdx.x == 0) {
        unsigned v;
        unsigned long long dead = t0 + 2000000000ull;
        do {
            asm volatile("ld.acquire.gpu.u32 %0, [%1];" : "=r"(v) : "l"(flags + blockIdx.x) : "memory");
            if (gtime() > dead) { __trap(); }
        } while ((int)v < launch - 1);
    }
    __syncthreads();
    const unsigned long long t1 = gtime();
    const unsigned long long dur = base_ns + (unsigned long long)blockIdx.x * step_ns;
    smem[threadIdx.x] = (float)launch;
    while (gtime() - t1 < dur) { }
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned smid;
        asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
        unsigned long long *r = times + ((size_t)launch * gridDim.x + blockIdx.x) * 4;
        r[0] = t0; r[1] = t1; r[2] = gtime(); r[3] = smid;
        asm volatile("st.release.gpu.u32 [%0], %1;" :: "l"(flags + blockIdx.x), "r"((unsigned)launch) : "memory");
    }
}

enum Between { kNone, kRecord, kWaitDone, kH2D, kD2D };

static bool run(const char *name, int ncta, bool pdl, bool coop, Between between, bool two_streams)
{
    const int L = 5, smem = 200 * 1024;
    const unsigned base = 2000000, step = 20000;
    printf("%s\n", name);
    CK(cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    unsigned long long *times; unsigned *flags; float *a, *b, *pin;
    CK(cudaMalloc(&times, sizeof(unsigned long long) * 4 * ncta * (L + 1)));
    CK(cudaMalloc(&flags, sizeof(unsigned) * ncta));
    CK(cudaMalloc(&a, 1 << 20)); CK(cudaMalloc(&b, 1 << 20)); CK(cudaMallocHost(&pin, 4096));
    CK(cudaMemset(flags, 0, sizeof(unsigned) * ncta));
    cudaStream_t st[2], other; CK(cudaStreamCreate(&st[0])); CK(cudaStreamCreate(&st[1])); CK(cudaStreamCreate(&other));
    cudaEvent_t ev, done_ev, e0, e1;
    CK(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming)); CK(cudaEventCreateWithFlags(&done_ev, cudaEventDisableTiming));
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    CK(cudaMemsetAsync(a, 0, 64, other)); CK(cudaEventRecord(done_ev, other)); CK(cudaStreamSynchronize(other));
    CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(e0, st[0]));
    for (int l = 1; l <= L; ++l) {
        cudaStream_t s = st[two_streams ? (l & 1) : 0];
        cudaLaunchConfig_t lc = {};
        lc.gridDim = dim3(ncta); lc.blockDim = dim3(320); lc.dynamicSmemBytes = smem; lc.stream = s;
        cudaLaunchAttribute at[2]; int na = 0;
        if (pdl) { at[na].id = cudaLaunchAttributeProgrammaticStreamSerialization; at[na].val.programmaticStreamSerializationAllowed = 1; ++na; }
        if (coop) { at[na].id = cudaLaunchAttributeCooperative; at[na].val.cooperative = 1; ++na; }
        lc.attrs = at; lc.numAttrs = na;
        int trig = pdl ? 1 : 0;
        CK(cudaLaunchKernelEx(&lc, probe_kernel, times, flags, l, L, trig, base, step));
        if (l < L) {
            if (between == kRecord) CK(cudaEventRecord(ev, s));
            if (between == kWaitDone) CK(cudaStreamWaitEvent(s, done_ev, 0));
            if (between == kH2D) CK(cudaMemcpyAsync(a, pin, 1024, cudaMemcpyHostToDevice, s));
            if (between == kD2D) CK(cudaMemcpyAsync(b, a, 1 << 20, cudaMemcpyDeviceToDevice, s));
        }
    }
    if (two_streams) { CK(cudaStreamSynchronize(st[1])); }
    CK(cudaEventRecord(e1, st[0]));
    CK(cudaDeviceSynchronize());
    float ms = 0; CK(cudaEventElapsedTime(&ms, e0, e1));
    std::vector<unsigned long long> h((size_t)4 * ncta * (L + 1));
    CK(cudaMemcpy(h.data(), times, h.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
    unsigned long long origin = ~0ull;
    for (int c = 0; c < ncta; ++c) origin = std::min(origin, h[((size_t)1 * ncta + c) * 4]);
    for (int l = 1; l <= L; ++l) {
        unsigned long long s0 = ~0ull, smax = 0, emin = ~0ull, emax = 0, w0 = 0;
        for (int c = 0; c < ncta; ++c) {
            const unsigned long long *r = &h[((size_t)l * ncta + c) * 4];
            s0 = std::min(s0, r[0]); smax = std::max(smax, r[0]); emin = std::min(emin, r[2]); emax = std::max(emax, r[2]);
            if (c == 0) w0 = r[1];
        }
        printf("  launch %d: first CTA in %.3f  last CTA in %.3f  CTA0 past its wait %.3f  first out %.3f  last out %.3f ms\n",
               l, (s0 - origin) * 1e-6, (smax - origin) * 1e-6, (w0 - origin) * 1e-6, (emin - origin) * 1e-6, (emax - origin) * 1e-6);
    }
    printf("  %d launches: %.3f ms (serial: %.1f, fully overlapped: %.1f)\n", L, ms,
           L * (base + (ncta - 1) * step) * 1e-6, (L * base + (ncta - 1) * step) * 1e-6);
    cudaFree(times); cudaFree(flags); cudaFree(a); cudaFree(b); cudaFreeHost(pin);
    cudaStreamDestroy(st[0]); cudaStreamDestroy(st[1]); cudaStreamDestroy(other);
    return true;
}

int main(int argc, char **argv)
{
    int ncta = argc > 1 ? atoi(argv[1]) : 129;
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    printf("%s, %d SMs, %d CTAs of 320 threads + 200 KB\n", p.name, p.multiProcessorCount, ncta);
    run("stream order", ncta, false, false, kNone, false);
    run("PDL", ncta, true, false, kNone, false);
    run("PDL + cooperative", ncta, true, true, kNone, false);
    run("PDL, event record between", ncta, true, false, kRecord, false);
    run("PDL, wait on a finished event between", ncta, true, false, kWaitDone, false);
    run("PDL, 1 KB H2D copy between", ncta, true, false, kH2D, false);
    run("PDL, 1 MB D2D copy between", ncta, true, false, kD2D, false);
    run("PDL, 148 CTAs", 148, true, false, kNone, false);
    run("two streams, no PDL", ncta, false, false, kNone, true);
    return 0;
}
