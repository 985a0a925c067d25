/* hw2: bubble sort of n floats, printed as "%.6e " each then a newline
 * (ref: hw2/src/main.c:4-15,34-37).  This version stops as soon as a sweep makes
 * no swap and shrinks the sweep to the last swap position. */
#include <stdio.h>
#include <stdlib.h>

static void bubble_sort(float* v, int n) {
  int bound = n - 1;
  while (bound > 0) {
    int last_swap = 0;
    for (int j = 0; j < bound; ++j) {
      if (v[j] > v[j + 1]) {
        float t = v[j];
        v[j] = v[j + 1];
        v[j + 1] = t;
        last_swap = j;
      }
    }
    bound = last_swap;
  }
}

int main(void) {
  int n;
  if (scanf("%d", &n) != 1 || n < 0) return 1;
  float* v = (float*)malloc(sizeof(float) * (size_t)(n > 0 ? n : 1));
  if (!v) {
    fprintf(stderr, "out of memory\n");
    return 1;
  }
  for (int i = 0; i < n; ++i)
    if (scanf("%f", &v[i]) != 1) return 1;
  bubble_sort(v, n);
  for (int i = 0; i < n; ++i) printf("%.6e ", v[i]);
  putchar('\n');
  free(v);
  return 0;
}
