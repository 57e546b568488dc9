/*
 * b200_install_shim.c -- TEST INFRASTRUCTURE: loads the generated b200_rtcd_install() next to an
 * already built reference library (oracle/_ref/libaomref_avx2.so, run-time dispatch) so that the
 * UNMODIFIED aomenc_avx2 runs with the `_b200` flavour bound:
 *     LD_PRELOAD=oracle/_ref/libb200install.so B200_RTCD_GROUPS=<mask> oracle/_ref/aomenc_avx2 ...
 * The constructor runs the library's own RTCD initialisers first (they are once-guarded, so the
 * encoder's later calls do not overwrite the bindings).
 */
#include <stdio.h>
#include <stdlib.h>

void aom_dsp_rtcd(void);
void av1_rtcd(void);
void aom_scale_rtcd(void);
int b200_rtcd_install(unsigned groups);

int b200_install_shim_bound = -2; /* -2: not requested, -1: no device, >= 0: symbols bound */

__attribute__((constructor)) static void b200_install_shim_init(void) {
  const char *g = getenv("B200_RTCD_GROUPS");
  if (!g || !*g) return;
  aom_dsp_rtcd();
  av1_rtcd();
  aom_scale_rtcd();
  b200_install_shim_bound = b200_rtcd_install((unsigned)strtoul(g, NULL, 0));
  fprintf(stderr, "b200_rtcd_install(groups=%s): %d RTCD symbols bound to the _b200 flavour\n", g,
          b200_install_shim_bound);
}
