// CPU check of fluid_sim_b200/csrc/glibc_rand.h against this machine's libc rand().
#include "glibc_rand.h"
#include <cstdio>
#include <cstdlib>
#include <initializer_list>

int main()
{
    using namespace fsb;
    uint32_t w[kRandDeg];
    rand_seed_window(1, w);
    srand(1);
    for (int i = 0; i < 100000; i++)
        if ((int)window_next(w) != rand()) { printf("sequential mismatch at %d\n", i); return 1; }
    for (unsigned seed : {1u, 12345u})
        for (unsigned long long n : {1ull, 5ull, 31ull, 512ull, 16384ull, 123457ull, 3000000ull}) {
            uint32_t a[kRandDeg];
            rand_seed_window(seed, a);
            RandJump J = jump_matrix(n);
            apply_jump(J, a);
            srand(seed);
            for (unsigned long long i = 0; i < n; i++) rand();
            for (int i = 0; i < 64; i++)
                if ((int)window_next(a) != rand()) { printf("jump %llu mismatch at %d\n", n, i); return 1; }
        }
    printf("ok\n");
    return 0;
}
