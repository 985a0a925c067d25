/* lab3 stdin-parser smoke tool (ref: lab3/src/test_read_input.c): reads
 * "nc" then per class "np x y ..." and echoes what it understood. */
#include <stdio.h>
#include <stdlib.h>

int main(void) {
  int nc;
  if (scanf("%d", &nc) != 1 || nc < 0) {
    fprintf(stderr, "expected the class count\n");
    return 1;
  }
  for (int c = 0; c < nc; ++c) {
    int np;
    if (scanf("%d", &np) != 1 || np < 0) {
      fprintf(stderr, "class %d: expected the point count\n", c + 1);
      return 1;
    }
    printf("class %d: %d points\n", c + 1, np);
    for (int p = 0; p < np; ++p) {
      int x, y;
      if (scanf("%d %d", &x, &y) != 2) {
        fprintf(stderr, "class %d: point %d is missing\n", c + 1, p + 1);
        return 1;
      }
      printf("  (%d, %d)\n", x, y);
    }
  }
  return 0;
}
