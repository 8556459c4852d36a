/* Stand-in for the reference's missing src/zlib/zutil.h (submodule madler/zlib @5a82f71 is
 * absent from /root/reference).  TEST INFRASTRUCTURE ONLY: lets the
 * untouched reference sources src/inflate.c, src/inffast.c, src/inftrees.c compile in
 * place into oracle/_ref/.  Consumers: reference src/inflate.c:91-94, src/inftrees.c:7-8,
 * src/inffast.c:6-9.  Nothing here is reference code; it only names what those files use. */
#ifndef ZSB_ORACLE_ZUTIL_SHIM_H
#define ZSB_ORACLE_ZUTIL_SHIM_H
#include <zlib.h>      /* system zlib 1.3: z_stream, Z_* codes, crc32(), adler32() */
#include <stddef.h>
#include <stdlib.h>
#include <string.h>
#ifdef ZSB_ORACLE_TRACE
#  include <stdio.h>
#endif

#define local static
#define ZLIB_INTERNAL __attribute__((visibility("hidden")))
#ifndef DEF_WBITS
#  define DEF_WBITS MAX_WBITS
#endif
#define zmemcpy memcpy
#define zmemzero(dest, len) memset(dest, 0, len)

#ifdef ZSB_ORACLE_TRACE
extern int z_verbose;
#  define Trace(x)   do { if (z_verbose >= 0) fprintf x; } while (0)
#  define Tracev(x)  do { if (z_verbose > 0) fprintf x; } while (0)
#  define Tracevv(x) do { if (z_verbose > 1) fprintf x; } while (0)
#else
#  define Trace(x)
#  define Tracev(x)
#  define Tracevv(x)
#endif

#define ZALLOC(strm, items, size) (*((strm)->zalloc))((strm)->opaque, (items), (size))
#define ZFREE(strm, addr) (*((strm)->zfree))((strm)->opaque, (voidpf)(addr))
#define ZSWAP32(q) ((((q) >> 24) & 0xff) + (((q) >> 8) & 0xff00) + \
                    (((q) & 0xff00) << 8) + (((q) & 0xff) << 24))

voidpf ZLIB_INTERNAL zcalloc(voidpf opaque, unsigned items, unsigned size);
void ZLIB_INTERNAL zcfree(voidpf opaque, voidpf ptr);
#endif
