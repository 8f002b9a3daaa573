/* include/lp_b200.h must be usable from plain C (the boundary is a C ABI): compiled with gcc -std=c99 -pedantic. */
#include "lp_b200.h"

int lpb200_header_probe(void)
{
    lpb200_comm_t *comm = 0;
    (void)comm;
    return LPB200_STATUS_BOUNDED + LPB200_RULE_DEVEX + LPB200_DTYPE_I8 + LPB200_PIVOT_PARTIAL + (int)sizeof(int32_t);
}
