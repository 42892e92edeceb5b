/* config.h for compiling the vendored htslib 1.23 from /root/reference with plain gcc
 * (no ./configure run; see oracle/Makefile).  zlib only: no bz2/lzma/curl/libdeflate,
 * so BGZF inflate + CRC go through system zlib exactly as SURVEY.md's probe table found. */
#ifndef _XOPEN_SOURCE
#define _XOPEN_SOURCE 600
#endif
#define HAVE_LIBZ 1
#define HAVE_DRAND48 1
#define HAVE_FSYNC 1
#define HAVE_FDATASYNC 1
#define HAVE_GMTIME_R 1
#define HAVE_MMAP 1
#define HAVE_GETPAGESIZE 1
#define HAVE_CLOCK_GETTIME_CPUTIME 1
#define HAVE_STDINT_H 1
#define HAVE_INTTYPES_H 1
#define HAVE_STDLIB_H 1
#define HAVE_STRING_H 1
#define HAVE_UNISTD_H 1
