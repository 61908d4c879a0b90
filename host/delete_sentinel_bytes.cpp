// delete_sentinel_bytes -- the reference's helper (tools/delete-sentinel-bytes/src/main.cpp:45-104) with the
// filter on the GPU: erase every byte of value 255 from FILE and write the result on standard output.
// Same usage text, same exit codes, same progress / summary lines on stderr.  The file is streamed through
// two pinned buffers; every chunk is compacted on the device (psc_delete_sentinel_bytes_d) and chunks are
// written in order, so the result is the stable filter of the whole file.
#include <sys/time.h>
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include "../include/psascan_cuda.h"

static long double wallclock() {
  timeval tim;
  gettimeofday(&tim, NULL);
  return tim.tv_sec + (tim.tv_usec / 1000000.L);
}

static void die(const char *what) {
  std::fprintf(stderr, "\nError: %s: %s\n", what, psc_last_error());
  std::exit(EXIT_FAILURE);
}

int main(int argc, char **argv) {
  if (argc != 2) {
    std::fprintf(stderr, "Usage: %s FILE\nErase all bytes with value 255 "
        "from FILE. Write result on standard output.\n", argv[0]);
    std::exit(EXIT_FAILURE);
  }
  std::FILE *f = std::fopen(argv[1], "r");
  if (f == NULL) {
    std::perror(argv[1]);
    std::exit(EXIT_FAILURE);
  }
  std::fseek(f, 0L, SEEK_END);
  long size = std::ftell(f);
  std::rewind(f);

  psc_context *ctx = NULL;
  if (psc_init(-1, 0, &ctx) != PSC_OK) die("cannot initialise the GPU");
  const size_t bufsize = (size_t)64 << 20;
  void *pin_in = NULL, *pin_out = NULL, *in_d = NULL, *out_d = NULL;
  if (psc_pinned_alloc(ctx, bufsize, &pin_in) != PSC_OK || psc_pinned_alloc(ctx, bufsize, &pin_out) != PSC_OK) die("pinned buffers");
  if (psc_device_alloc(ctx, bufsize, &in_d) != PSC_OK || psc_device_alloc(ctx, bufsize, &out_d) != PSC_OK) die("device buffers");

  long double start = wallclock();
  std::size_t elems, count = 0, total = 0;
  while ((elems = std::fread(pin_in, 1, bufsize, f)) > 0) {
    total += elems;
    count += elems;
    long kept = 0;
    if (psc_copy_h2d(ctx, in_d, pin_in, elems) != PSC_OK) die("host to device copy");
    if (psc_delete_sentinel_bytes_d(ctx, (const uint8_t *)in_d, (long)elems, (uint8_t *)out_d, &kept) != PSC_OK) die("filter");
    if (kept > 0) {
      if (psc_copy_d2h(ctx, pin_out, out_d, (size_t)kept) != PSC_OK) die("device to host copy");
      std::fwrite(pin_out, 1, (size_t)kept, stdout);
    }
    if (count > (64L << 20)) {
      count = 0;
      long double elapsed = wallclock() - start;
      long double mib = (long double)total / (1L << 20);
      std::fprintf(stderr, "Processed %.0LfMiB (%.1Lf%%). Speed: %.2LfMiB/s.\r",
          mib, (100.L * total) / size, mib / elapsed);
    }
  }
  std::fclose(f);
  psc_pinned_free(ctx, pin_in);
  psc_pinned_free(ctx, pin_out);
  psc_device_free(ctx, in_d);
  psc_device_free(ctx, out_d);
  psc_shutdown(ctx);

  long double elapsed = wallclock() - start;
  long double mib = (long double)size / (1L << 20);
  std::fprintf(stderr, "Processed %.0LfMiB (100.0%%). Speed: %.2LfMiB/s.\n",
      mib, mib / elapsed);
}
