/* TEST INFRASTRUCTURE ONLY (see oracle/README.md).
 * Replacement for the five malloc_count_* symbols bfq_ext.cpp:1075-1110 prints.
 * The reference's malloc_count wraps malloc/free only and aborts with gcc-13's
 * libgomp (aligned_alloc + free) -- SURVEY.md F5.  This stub changes no algorithm. */
#include <stddef.h>
size_t malloc_count_current(void)   { return 0; }
size_t malloc_count_peak(void)      { return 0; }
size_t malloc_count_peak_curr(void) { return 0; }
void   malloc_count_reset_peak(void){ }
void   malloc_count_print_status(void){ }
