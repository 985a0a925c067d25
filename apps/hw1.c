/* hw1: roots of a*x^2 + b*x + c = 0 in single precision.
 * Output contract (ref: hw1/src/main.c:8-30): "any" / "incorrect" for the
 * degenerate cases, one "%.6f" root for a linear or double root, two roots
 * "(-b+sqrt(D))/2a (-b-sqrt(D))/2a" for D > 0, "imaginary" for D < 0. */
#include <math.h>
#include <stdio.h>

int main(void) {
  float a, b, c;
  if (scanf("%f %f %f", &a, &b, &c) != 3) return 1;
  if (a == 0.0f && b == 0.0f) {
    puts(c == 0.0f ? "any" : "incorrect");
    return 0;
  }
  if (a == 0.0f) {
    printf("%.6f\n", -c / b);
    return 0;
  }
  const float disc = b * b - 4 * a * c;
  const float two_a = 2 * a;
  if (disc > 0.0f) {
    const float root = sqrtf(disc);
    printf("%.6f %.6f\n", (-b + root) / two_a, (-b - root) / two_a);
  } else if (disc == 0.0f) {
    printf("%.6f\n", -b / two_a);
  } else {
    puts("imaginary");
  }
  return 0;
}
