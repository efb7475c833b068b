/* metis_probe.c -- width of idx_t in the METIS build that ships with the CUDA toolkit (libmetis_static.a has no header here).
 * METIS_SetDefaultOptions fills METIS_NOPTIONS = 40 entries of idx_t with -1: the bytes it changes in a poisoned buffer are
 * 40 * sizeof(idx_t).  Prints "idx_t bytes: 4" or "idx_t bytes: 8".
 *   gcc -O1 -o /tmp/metis_probe tools/metis_probe.c /usr/local/cuda/targets/x86_64-linux/lib/libmetis_static.a -lm
 */
#include <stdio.h>
#include <string.h>

int METIS_SetDefaultOptions(void *options);

int main(void) {
  unsigned char buf[1024];
  memset(buf, 0x5a, sizeof buf);
  METIS_SetDefaultOptions(buf);
  int changed = 0;
  for (int i = 0; i < (int)sizeof buf; ++i) changed += buf[i] != 0x5a;
  printf("bytes changed: %d\nidx_t bytes: %d\n", changed, changed / 40);
  return (changed == 160 || changed == 320) ? 0 : 1;
}
