/* Stand-in for the reference's missing src/zlib/inffast.h: one prototype, consumed by
 * reference src/inflate.c:94 and src/inffast.c:9.  TEST INFRASTRUCTURE ONLY. */
#include <zlib.h>
void __attribute__((visibility("hidden"))) inflate_fast(z_streamp strm, unsigned start);
