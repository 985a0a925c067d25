// lab1 contract binary: c = a - b over doubles.
//   cuda_exe          stdin "n a.. b.."            stdout "%.10e "*n      (ref: lab1/src/main.cu)
//   cuda_exe_to_plot  stdin "grid block n a.. b.." stdout time line + vector (ref: lab1/src/to_plot.cu)
// The to_plot build launches <<<grid, block>>> exactly as given (the harness sweeps
// launch geometry); grid or block <= 0 selects the tuned B200 kernel.
#include "b200labs.h"
#include "cli_common.h"

int main() {
  cli::Input in;
  in.slurp(stdin);
  int grid = 0, block = 0, n = 0;
#ifdef TO_PLOT
  if (!in.next_int(grid) || !in.next_int(block)) cli::die("expected launch dims: grid block");
#endif
  if (!in.next_int(n) || n < 0) cli::die("expected vector size n");
  std::vector<double> a(n), b(n), c(n);
  for (int i = 0; i < n; ++i)
    if (!in.next_f64(a[i])) cli::die("first vector is shorter than n");
  for (int i = 0; i < n; ++i)
    if (!in.next_f64(b[i])) cli::die("second vector is shorter than n");

  double *da = nullptr, *db = nullptr, *dc = nullptr;
  const size_t bytes = sizeof(double) * (size_t)n;
  CSC(cudaMalloc(&da, bytes ? bytes : 8));
  CSC(cudaMalloc(&db, bytes ? bytes : 8));
  CSC(cudaMalloc(&dc, bytes ? bytes : 8));
  CSC(cudaMemcpy(da, a.data(), bytes, cudaMemcpyHostToDevice));
  CSC(cudaMemcpy(db, b.data(), bytes, cudaMemcpyHostToDevice));

  LIBOK(b200_preload_lab1());
  float ms = cli::time_device_ms([&] {
    if (grid > 0 && block > 0)
      LIBOK(b200_vecsub_f64_geom(da, db, dc, (size_t)n, grid, block, nullptr));
    else
      LIBOK(b200_vecsub_f64(da, db, dc, (size_t)n, nullptr));
  });
  CSC(cudaMemcpy(c.data(), dc, bytes, cudaMemcpyDeviceToHost));
#ifdef TO_PLOT
  printf("CUDA execution time: <%f ms>\n", ms);
#else
  (void)ms;
#endif
  {
    cli::Out out(stdout);
    for (int i = 0; i < n; ++i) out.sci(c[i], 10);
  }
  cudaFree(da), cudaFree(db), cudaFree(dc);
  return 0;
}
