/* Default allocator the reference expects from zutil.c (missing submodule).
 * TEST INFRASTRUCTURE ONLY (oracle/_ref build). */
#include "zutil.h"
#ifdef ZSB_ORACLE_TRACE
int z_verbose = 0;
#endif
voidpf ZLIB_INTERNAL zcalloc(voidpf opaque, unsigned items, unsigned size) {
    (void)opaque;
    return calloc(items, size);
}
void ZLIB_INTERNAL zcfree(voidpf opaque, voidpf ptr) {
    (void)opaque;
    free(ptr);
}
