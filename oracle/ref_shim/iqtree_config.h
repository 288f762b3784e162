/* Hand-written stand-in for the header CMake generates from iqtree_config.h.in (test infrastructure:
 * lets oracle/Makefile.ref compile the reference's likelihood translation units in place). */
#define iqtree_VERSION_MAJOR 2
#define iqtree_VERSION_MINOR 2
#define iqtree_VERSION_PATCH "0"
#define HAVE_GETTIMEOFDAY
#define HAVE_GETRUSAGE
#define HAVE_STRNDUP
#define HAVE_STRTOK_R
