// lab1 CPU binary: same stdin as cuda_exe (no launch-dims prefix), prints
// "CPU execution time: <t ms>" then the vector (ref: lab1/src/main.c:54-63).
// Built with -DNO_TIME_LINE it is the submission-style cpu_exe (vector only).
#include <chrono>

#include "b200labs.h"
#include "cli_common.h"

int main() {
  cli::Input in;
  in.slurp(stdin);
  int n = 0;
  if (!in.next_int(n) || n < 0) cli::die("expected vector size n");
  std::vector<double> a(n), b(n), c(n);
  for (int i = 0; i < n; ++i)
    if (!in.next_f64(a[i])) cli::die("first vector is shorter than n");
  for (int i = 0; i < n; ++i)
    if (!in.next_f64(b[i])) cli::die("second vector is shorter than n");
  auto t0 = std::chrono::steady_clock::now();
  b200_cpu_vecsub_f64(a.data(), b.data(), c.data(), (size_t)n);
  auto t1 = std::chrono::steady_clock::now();
#ifndef NO_TIME_LINE
  printf("CPU execution time: <%f ms>\n", std::chrono::duration<double, std::milli>(t1 - t0).count());
#else
  (void)t0, (void)t1;
#endif
  cli::Out out(stdout);
  for (int i = 0; i < n; ++i) out.sci(c[i], 10);
  return 0;
}
