/* Hand-written stand-in for the header the reference's CMake generates from
 * iqtree_config.h.in (CMakeLists.txt:323-326).  TEST INFRASTRUCTURE for oracle/_ref only. */
#define iqtree_VERSION_MAJOR 1
#define iqtree_VERSION_MINOR 4
#define iqtree_VERSION_PATCH "3-beta"
#define HAVE_GETTIMEOFDAY
#define HAVE_GETRUSAGE
