// lab2 contract binary: Roberts-cross filter over "int w, int h, RGBA[]" files.
//   cuda_exe          stdin "in out"              no stdout            (ref: lab2/src/main.cu)
//   to_plot_cuda_exe  stdin "bx by gx gy in out"  time line + FINISHED (ref: lab2/src/to_plot.cu)
// to_plot honours the block/grid it is given (launch-geometry sweep); any dim
// <= 0 selects the tuned kernel.  B200LABS_ROBERTS=stream|tma|cluster|generic forces a variant.
// The kernel is timed cold, one launch, like the reference (B200LABS_WARM=1: after a warm-up).
// Large images go through the pipelined file <-> GPU path of cli_pipeline.h; the image is
// never held in host memory as a whole, so 64-bit sizes (w * h >= 2^31 pixels) work.
#include "b200labs.h"
#include "cli_pipeline.h"

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
  const bool geom = bx > 0 && by > 0 && gx > 0 && gy > 0;
  const int variant = variant_from_env();

  const int in_fd = open(in_path.c_str(), O_RDONLY);
  if (in_fd < 0) cli::die("cannot open input image '%s'", in_path.c_str());
  int wh[2] = {0, 0};
  if (pread(in_fd, wh, 8, 0) != 8) cli::die("'%s': no w/h header", in_path.c_str());
  const int w = wh[0], h = wh[1];
  if (w <= 0 || h <= 0) cli::die("'%s': non-positive image size", in_path.c_str());
  const size_t bytes = (size_t)w * (size_t)h * 4;
  const bool piped = bytes >= cli::kPipeMin && !cli::same_file(in_path, out_path);

  cli::OutputFile out;
  if (piped) out.begin(out_path, wh, 8, 8 + bytes);  // pages get allocated while CUDA starts up
  void *d_in = nullptr, *d_out = nullptr;
  CSC(cudaMalloc(&d_in, bytes));
  CSC(cudaMalloc(&d_out, bytes));
  LIBOK(b200_preload_lab2());
  auto run = [&] {
    return cli::time_device_ms([&] {
      if (geom)
        LIBOK(b200_roberts_geom(d_in, d_out, w, h, bx, by, gx, gy, nullptr));
      else
        LIBOK(b200_roberts(d_in, d_out, w, h, nullptr, variant, nullptr));
    });
  };
  [[maybe_unused]] float ms = 0;  // printed by the to_plot build only
  if (piped) {
    cli::PinnedRing ring;
    cli::file_to_device(in_fd, 8, d_in, bytes, ring, in_path.c_str());
    close(in_fd);
    ms = run();
#ifdef TO_PLOT
    printf("CUDA execution time: <%f ms>\n", ms);
#endif
    out.ready();
    cli::device_to_file(out.fd, 8, d_out, bytes, ring, out_path.c_str());
    out.finish();
  } else {
    std::vector<uint32_t> px((size_t)w * h);
    cli::pread_full(in_fd, px.data(), bytes, 8, in_path.c_str());
    close(in_fd);
    CSC(cudaMemcpy(d_in, px.data(), bytes, cudaMemcpyHostToDevice));
    ms = run();
    CSC(cudaMemcpy(px.data(), d_out, bytes, cudaMemcpyDeviceToHost));
#ifdef TO_PLOT
    printf("CUDA execution time: <%f ms>\n", ms);
#endif
    cli::write_image(out_path, w, h, px.data());
  }
  cudaFree(d_in), cudaFree(d_out);
#ifdef TO_PLOT
  printf("FINISHED!\n");
#endif
  return 0;
}
