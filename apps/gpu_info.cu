// gpu_info: the eight labelled lines of ref: gpu_info/src/main.cu:9-16 for
// device 0; `gpu_info --extended` appends B200-relevant extras (L2, opt-in shared
// memory, cluster launch, peer devices); `gpu_info N` queries device N.
#include "b200labs.h"
#include "cli_common.h"

int main(int argc, char** argv) {
  int dev = 0, ext = 0;
  for (int i = 1; i < argc; ++i) {
    if (!strcmp(argv[i], "--extended")) ext = 1;
    else dev = atoi(argv[i]);
  }
  char buf[2048];
  int n = b200_gpu_info(dev, buf, sizeof buf, ext);
  if (n <= 0 || b200_last_error()[0]) {
    fprintf(stderr, "[ERROR CUDA] %s\n", b200_last_error());
    return 1;
  }
  fputs(buf, stdout);
  return 0;
}
