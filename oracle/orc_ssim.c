/* oracle/orc_ssim.c — TEST INFRASTRUCTURE. Placeholder translation unit; the D-SSIM term
 * (a "next" row) is added here when the CUDA loss kernel lands. */
#include "orc_common.h"
int orc_ssim_available(void) { return 0; }
