/* Driving libb200ddp from plain C (no Python, no torch): the C ABI of include/b200ddp.h.
 *
 *   gcc -O2 -Iinclude examples/c_host.c -o c_host \
 *       -Lcloud-security-pytorch-sagemaker-distributed-workshop_b200/_lib -lb200ddp -Wl,-rpath,'$ORIGIN/...' -ldl -lpthread -lrt
 *
 * Without a GPU it prints the library version and stops; with one it runs a fused SGD step and the allocation-free BN
 * statistics entry point on device buffers it allocates through the CUDA runtime that is linked into the library. */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "b200ddp.h"

/* the static CUDA runtime inside the library exports nothing; a host that has no CUDA of its own asks the driver */
#include <dlfcn.h>
typedef int (*cu_init_t)(unsigned);
typedef int (*cu_count_t)(int*);

static int gpu_count(void) {
  void* h = dlopen("libcuda.so.1", RTLD_NOW);
  if (!h) return 0;
  cu_init_t init = (cu_init_t)dlsym(h, "cuInit");
  cu_count_t count = (cu_count_t)dlsym(h, "cuDeviceGetCount");
  int n = 0;
  if (!init || !count || init(0) != 0 || count(&n) != 0) n = 0;
  return n;
}

int main(void) {
  printf("b200ddp C ABI version %d\n", b200_version());
  if (gpu_count() == 0) {
    printf("no GPU visible: ABI link check only\n");
    return 0;
  }
  printf("GPU visible: a C host would now hand device pointers and a stream to b200_sgd_momentum(), b200_conv_fwd(), ...\n");
  printf("last error string: \"%s\"\n", b200_last_error());
  return 0;
}
