/* strm->msg strings of inflate(); see oracle_api.h.  TEST INFRASTRUCTURE ONLY. */
#include "oracle_api.h"
const char *const orc_messages[ORC_NUM_MESSAGES] = {
    0,
    "incorrect header check",               /* 1  inflate.c:652 */
    "unknown compression method",           /* 2  inflate.c:657,682 */
    "invalid window size",                  /* 3  inflate.c:666 */
    "unknown header flags set",             /* 4  inflate.c:687 */
    "header crc mismatch",                  /* 5  inflate.c:803 */
    "invalid block type",                   /* 6  inflate.c:865 */
    "invalid stored block lengths",         /* 7  inflate.c:874 */
    "too many length or distance symbols",  /* 8  inflate.c:916 */
    "invalid code lengths set",             /* 9  inflate.c:940 */
    "invalid bit length repeat",            /* 10 inflate.c:964,987 */
    "invalid code -- missing end-of-block", /* 11 inflate.c:1001 */
    "invalid literal/lengths set",          /* 12 inflate.c:1016 */
    "invalid distances set",                /* 13 inflate.c:1026 */
    "invalid literal/length code",          /* 14 inflate.c:1080, inffast.c:283 */
    "invalid distance code",                /* 15 inflate.c:1118, inffast.c:268 */
    "invalid distance too far back",        /* 16 inflate.c:1150, inffast.c:172 */
    "incorrect data check",                 /* 17 inflate.c:1209 */
    "incorrect length check",               /* 18 inflate.c:1223 */
};
