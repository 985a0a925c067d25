// lab2 contract binary: Roberts-cross filter over "int w, int h, RGBA[]" files.
//   cuda_exe          stdin "in out"              no stdout            (ref: lab2/src/main.cu)
//   to_plot_cuda_exe  stdin "bx by gx gy in out"  time line + FINISHED (ref: lab2/src/to_plot.cu)
// to_plot honours the block/grid it is given (launch-geometry sweep); any dim
// <= 0 selects the tuned kernel.  B200LABS_ROBERTS=stream|tma|cluster|generic forces a variant.
#include "b200labs.h"
#include "cli_common.h"

static int variant_from_env() {
  const char* v = getenv("B200LABS_ROBERTS");
  if (!v) return B200_ROBERTS_AUTO;
  if (!strcmp(v, "stream")) return B200_ROBERTS_STREAM;
  if (!strcmp(v, "tma")) return B200_ROBERTS_TMA;
  if (!strcmp(v, "generic")) return B200_ROBERTS_GENERIC;
  if (!strcmp(v, "cluster")) return B200_ROBERTS_CLUSTER;
  return B200_ROBERTS_AUTO;
}

int main() {
  cli::Input in;
  in.slurp(stdin);
  int bx = 0, by = 0, gx = 0, gy = 0;
#ifdef TO_PLOT
  if (!in.next_int(bx) || !in.next_int(by) || !in.next_int(gx) || !in.next_int(gy))
    cli::die("expected launch dims: block_x block_y grid_x grid_y");
#endif
  std::string in_path, out_path;
  if (!in.next_word(in_path) || !in.next_word(out_path)) cli::die("expected input and output paths");
  cli::Image im = cli::read_image(in_path);
  const size_t bytes = im.px.size() * 4;
  void *d_in = nullptr, *d_out = nullptr;
  CSC(cudaMalloc(&d_in, bytes));
  CSC(cudaMalloc(&d_out, bytes));
  CSC(cudaMemcpy(d_in, im.px.data(), bytes, cudaMemcpyHostToDevice));
  const bool geom = bx > 0 && by > 0 && gx > 0 && gy > 0;
  const int variant = variant_from_env();
  float ms = cli::time_device_ms([&] {
    if (geom)
      LIBOK(b200_roberts_geom(d_in, d_out, im.w, im.h, bx, by, gx, gy, nullptr));
    else
      LIBOK(b200_roberts(d_in, d_out, im.w, im.h, nullptr, variant, nullptr));
  });
  CSC(cudaMemcpy(im.px.data(), d_out, bytes, cudaMemcpyDeviceToHost));
  cudaFree(d_in), cudaFree(d_out);
#ifdef TO_PLOT
  printf("CUDA execution time: <%f ms>\n", ms);
#else
  (void)ms;
#endif
  cli::write_image(out_path, im.w, im.h, im.px.data());
#ifdef TO_PLOT
  printf("FINISHED!\n");
#endif
  return 0;
}
