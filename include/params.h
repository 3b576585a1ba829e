/* params.h -- compile-time defaults of the B200 kNN engine.
 *
 * Same three names as the reference's params.h:3-6, so a consumer TU that sizes its result rows with
 * DEFAULT_NB_PLANES (test_knearests.cu:152,158) keeps compiling.  Unlike the reference this file is
 * include-guarded and every macro can be overridden with -D, and in this engine k is also a RUN-TIME
 * value (knc_prepare(..., k, ...) in knearests_c.h): DEFAULT_NB_PLANES is only what the seven
 * reference-compatible kn_* entry points use when the caller gives no k.
 */
#ifndef KNEARESTS_B200_PARAMS_H
#define KNEARESTS_B200_PARAMS_H

/* Threads per block of the reference query kernel (params.h:3).  Kept for source compatibility; the
 * B200 query kernel picks its own block size and does not read this. */
#ifndef POINTS_PER_BLOCK
#define POINTS_PER_BLOCK 32
#endif

/* k, the number of neighbours per point (params.h:4).  The library is built with the same default. */
#ifndef DEFAULT_NB_PLANES
#define DEFAULT_NB_PLANES 50
#endif

/* Logging switch (params.h:6).  The engine's own chatter is a run-time switch instead: it goes to
 * stderr when the environment variable KN_VERBOSE is set to a non-zero value. */
#ifndef IF_VERBOSE
#define IF_VERBOSE(x) x
#endif

#endif /* KNEARESTS_B200_PARAMS_H */
