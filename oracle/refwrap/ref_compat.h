/* TEST INFRASTRUCTURE ONLY — force-included (-include) when compiling the UNMODIFIED reference
 * sources with GCC >= 11: reference src/utils.h:51-67 defines its own __rdtsc(), which clashes
 * with the compiler's <ia32intrin.h>.  Pulling the intrinsics header in first and then renaming
 * the identifier makes utils.h define (and every caller use) an equivalent, differently named
 * inline function; no reference source is modified or copied. */
#include <x86intrin.h>
#define __rdtsc ref_compat_rdtsc
