// lab3 contract binary: Mahalanobis pixel classifier, class index -> alpha byte.
//   cuda_exe          stdin "in out nc {np x y ..}"                 (ref: lab3/src/main.cu)
//   to_plot_cuda_exe  stdin "blocks threads" + the above, prints the time line
//                     (ref: lab3/src/to_plot.cu:76-86,174)
// Class statistics: exact integer moments reduced on the GPU (default), or the
// reference's two-pass host loop in input order with B200LABS_LAB3_STATS=host.
// As in the reference only the classify pass is inside the timed region, one cold launch
// (B200LABS_WARM=1: after a warm-up).  Large images use the pipelined file <-> GPU path.
#include "cli_pipeline.h"
#include "lab3_common.h"

int main() {
  cli::Input in;
  in.slurp(stdin);
  int blocks = 0, threads = 0;
#ifdef TO_PLOT
  if (!in.next_int(blocks) || !in.next_int(threads)) cli::die("expected launch dims: blocks threads");
#endif
  Lab3Spec spec = read_lab3_spec(in);
  const int in_fd = open(spec.in_path.c_str(), O_RDONLY);
  if (in_fd < 0) cli::die("cannot open input image '%s'", spec.in_path.c_str());
  int wh[2] = {0, 0};
  if (pread(in_fd, wh, 8, 0) != 8) cli::die("'%s': no w/h header", spec.in_path.c_str());
  const int w = wh[0], h = wh[1];
  if (w <= 0 || h <= 0) cli::die("'%s': non-positive image size", spec.in_path.c_str());
  check_points(spec, w, h);
  const int64_t npix = (int64_t)w * h;
  const size_t bytes = (size_t)npix * 4;
  const char* mode = getenv("B200LABS_LAB3_STATS");
  const bool host_stats = mode && !strcmp(mode, "host");  // needs the whole image on the host
  const bool piped = bytes >= cli::kPipeMin && !host_stats && !cli::same_file(spec.in_path, spec.out_path);

  cli::OutputFile out;
  if (piped) out.begin(spec.out_path, wh, 8, 8 + bytes);
  void* d_img = nullptr;
  double* d_params = nullptr;
  CSC(cudaMalloc(&d_img, bytes));
  CSC(cudaMalloc(&d_params, sizeof(double) * B200_LAB3_MAX_CLASSES * B200_LAB3_PARAMS));
  LIBOK(b200_preload_lab3());
  std::vector<uint32_t> px;
  cli::PinnedRing* ring = nullptr;
  if (piped) {
    ring = new cli::PinnedRing();
    cli::file_to_device(in_fd, 8, d_img, bytes, *ring, spec.in_path.c_str());
  } else {
    px.resize((size_t)npix);
    cli::pread_full(in_fd, px.data(), bytes, 8, spec.in_path.c_str());
    CSC(cudaMemcpy(d_img, px.data(), bytes, cudaMemcpyHostToDevice));
  }
  close(in_fd);

  if (host_stats) {
    std::vector<double> params((size_t)spec.nc * B200_LAB3_PARAMS);
    b200_cpu_lab3_stats_ref(px.data(), w, spec.pts.data(), spec.class_off.data(), spec.nc, params.data());
    CSC(cudaMemcpy(d_params, params.data(), params.size() * sizeof(double), cudaMemcpyHostToDevice));
  } else {
    int *d_pts = nullptr, *d_off = nullptr;
    long long* d_mom = nullptr;
    const size_t mom_bytes = sizeof(long long) * spec.nc * B200_LAB3_MOMENTS;
    CSC(cudaMalloc(&d_pts, sizeof(int) * (spec.pts.size() + 2)));
    CSC(cudaMalloc(&d_off, sizeof(int) * spec.class_off.size()));
    CSC(cudaMalloc(&d_mom, mom_bytes));
    CSC(cudaMemcpy(d_pts, spec.pts.data(), sizeof(int) * spec.pts.size(), cudaMemcpyHostToDevice));
    CSC(cudaMemcpy(d_off, spec.class_off.data(), sizeof(int) * spec.class_off.size(),
                   cudaMemcpyHostToDevice));
    CSC(cudaMemset(d_mom, 0, mom_bytes));
    LIBOK(b200_lab3_moments(d_img, w, 0, h, d_pts, d_off, spec.nc, d_mom, nullptr));
    LIBOK(b200_lab3_finalize(d_mom, spec.nc, d_params, nullptr));
    CSC(cudaDeviceSynchronize());
    cudaFree(d_pts), cudaFree(d_off), cudaFree(d_mom);
  }

  const bool geom = blocks > 0 && threads > 0;
  size_t scratch_bytes = geom ? 0 : b200_lab3_classify_scratch_bytes(npix, B200_CLASSIFY_AUTO);
  void* d_scratch = nullptr;
  if (scratch_bytes) CSC(cudaMalloc(&d_scratch, scratch_bytes));
  // classification is in place and idempotent (it only rewrites the alpha byte, which the
  // distances ignore), so an optional warm-up run does not change the result
  float ms = cli::time_device_ms([&] {
    if (geom)
      LIBOK(b200_lab3_classify_geom(d_img, npix, spec.nc, d_params, blocks, threads, nullptr));
    else
      LIBOK(b200_lab3_classify(d_img, npix, spec.nc, d_params, B200_CLASSIFY_AUTO, d_scratch,
                               scratch_bytes, nullptr));
  });
  if (piped) {
    out.ready();
    cli::device_to_file(out.fd, 8, d_img, bytes, *ring, spec.out_path.c_str());
    out.finish();
    delete ring;
  } else {
    CSC(cudaMemcpy(px.data(), d_img, bytes, cudaMemcpyDeviceToHost));
    cli::write_image(spec.out_path, w, h, px.data());
  }
  cudaFree(d_img), cudaFree(d_params);
  if (d_scratch) cudaFree(d_scratch);
#ifdef TO_PLOT
  printf("CUDA execution time: <%f ms>\n", ms);
#else
  (void)ms;
#endif
  return 0;
}
