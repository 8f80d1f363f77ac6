/* Using libewdml.so from plain C (C99): the host-side pieces that need no GPU -- symmetric-heap sub-allocator, counters,
 * collective-engine sizing.  The kernels take the same kind of calls with device pointers and a CUDA stream.
 *
 *   gcc -std=c99 -Iinclude examples/c_abi_demo.c -o /tmp/c_abi_demo -Lewdml/_C -lewdml -Wl,-rpath,$PWD/ewdml/_C && /tmp/c_abi_demo
 */
#include <stdio.h>
#include "ewdml.h"
int main(void) {
    ewdml_symheap* h = NULL;
    if (ewdml_symheap_create(&h, 1 << 20, 256)) { printf("err %s\n", ewdml_last_error()); return 1; }
    long long a = ewdml_symheap_alloc(h, 1000), b = ewdml_symheap_alloc(h, 70000);
    printf("version %d offsets %lld %lld used %lld\n", ewdml_version(), a, b, (long long)ewdml_symheap_used(h));
    ewdml_symheap_destroy(h);
    ewdml_counter_add(EWDML_CTR_SEND_CALLS, 3);
    printf("counter %lld window %lld\n", (long long)ewdml_counter_get(EWDML_CTR_SEND_CALLS), (long long)ewdml_engine_window_bytes(8, 1 << 20));
    return 0;
}
