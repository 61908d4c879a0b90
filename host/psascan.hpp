// host/psascan.hpp -- the reference's public entry point on top of the CUDA C ABI.
//
//   void pSAscan(std::string input_filename, std::string output_filename, std::string gap_filename,
//                long ram_use, long max_threads, bool verbose);
//
// Same signature, argument meaning and error behaviour (message to stderr, std::exit(EXIT_FAILURE))
// as /root/reference/include/psascan.hpp:137-141 and :53-131.  The text is streamed from the file to
// the GPU through a pinned staging buffer, the whole block pipeline runs in HBM
// (psc_construct_sa_d), and the 40-bit suffix array is streamed back to OUTFILE, so host memory use
// is two staging buffers regardless of the text length.  ram_use bounds those buffers; the block
// size comes from the HBM budget (psc_max_block_size) instead of ram_use / 5.2 (psascan.hpp:91),
// or from the PSASCAN_MAX_BLOCK environment variable (testing).  With several visible GPUs (PSASCAN_GPUS
// caps the number) and a text of several blocks, one host thread per GPU joins the library's NCCL
// communicator (psc_dist_init) and the sharded schedule of csrc/dist.cu runs; every GPU's slice of the
// suffix array is then written to OUTFILE in rank order.  gap_filename and max_threads are accepted for
// compatibility: the reference only uses gap_filename for temporary v-byte gap files that it deletes
// again (merge.hpp:178-179), gap arrays never leave the devices here, and the work runs on the GPUs, so
// the reference's "not enough memory to start threads" check (psascan.hpp:73-86) and its RLIMIT_NOFILE
// check (:101-114, four files per block) have nothing to guard: two files are open at any time.
#ifndef PSASCAN_B200_HOST_PSASCAN_HPP
#define PSASCAN_B200_HOST_PSASCAN_HPP

#include <sys/stat.h>
#include <sys/time.h>
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <thread>
#include <vector>

#include "../include/psascan_cuda.h"

namespace psascan_b200 {

inline double wclock() {
  timeval tim;
  gettimeofday(&tim, NULL);
  return tim.tv_sec + tim.tv_usec / 1e6;
}

inline void die_psc(const char *what) {
  std::fprintf(stderr, "\nError: %s: %s\n", what, psc_last_error());
  std::exit(EXIT_FAILURE);
}

inline long file_size_or_die(const std::string &fname) {
  struct stat st;
  if (stat(fname.c_str(), &st) != 0) {
    std::perror(fname.c_str());  // utils.cpp:60-63 style: report and exit
    std::exit(EXIT_FAILURE);
  }
  return (long)st.st_size;
}

// One GPU of the run: its context, the device copy of the text and (multi-GPU) its slice of the output.
struct GpuRank {
  psc_context *ctx = NULL;
  void *text_d = NULL, *out_d = NULL;
  long slice_beg = 0, slice_len = 0;
  int rc = 0;
  std::string err;
};

inline void pSAscan_impl(const std::string &input_filename, const std::string &output_filename,
                         const std::string &gap_filename, long ram_use, long max_threads, bool verbose) {
  if (ram_use < 6L) {  // psascan.hpp:57-60
    std::fprintf(stderr, "Error: not enough memory to run pSAscan.\n");
    std::exit(EXIT_FAILURE);
  }
  const long n = file_size_or_die(input_filename);
  if (n == 0) {  // the reference cannot handle it either (gap_array.hpp:57-60)
    std::fprintf(stderr, "\nError: the input file is empty\n");
    std::exit(EXIT_FAILURE);
  }
  const double start = wclock();

  // GPUs: all visible ones (PSASCAN_GPUS caps it); one rank per GPU, one host thread per rank
  int n_gpus = psc_device_count();
  if (n_gpus <= 0) {
    psc_context *probe = NULL;
    psc_init(-1, 0, &probe);  // fills psc_last_error with the reason
    die_psc("cannot initialise the GPU");
  }
  if (const char *e = std::getenv("PSASCAN_GPUS")) n_gpus = std::max(1, std::min(n_gpus, std::atoi(e)));
  std::vector<GpuRank> gpu(n_gpus);
  if (psc_init(n_gpus > 1 ? 0 : -1, 0, &gpu[0].ctx) != PSC_OK) die_psc("cannot initialise the GPU");
  char gpu_name[128] = "";
  size_t free_b = 0, total_b = 0;
  psc_device_info(gpu[0].ctx, gpu_name, sizeof(gpu_name), &free_b, &total_b);

  long max_block = psc_max_block_size(gpu[0].ctx, n);
  // External-memory mode: partial suffix arrays and gap arrays as files in the reference's formats, streaming merge.
  // Taken when the resident pipeline does not fit the HBM (text + 8 bytes per suffix + the output) or on request.
  bool em_mode = std::getenv("PSASCAN_EM") != NULL && std::atoi(std::getenv("PSASCAN_EM")) != 0;
  if (max_block <= 0 && !em_mode) em_mode = psc_max_block_size_files(gpu[0].ctx, n) > 0;
  if (em_mode) {
    max_block = psc_max_block_size_files(gpu[0].ctx, n);
    n_gpus = 1;
  }
  if (const char *e = std::getenv("PSASCAN_MAX_BLOCK")) max_block = std::max(1L, std::atol(e));
  if (max_block <= 0) {
    std::fprintf(stderr, "\nError: text of %ld bytes does not fit the HBM budget (%.1f GiB free)\n", n,
                 free_b / 1073741824.0);
    std::exit(EXIT_FAILURE);
  }
  // several GPUs only pay off with at least one block per GPU: a text that fits one block stays on one GPU
  // unless the block size was forced
  if (n_gpus > 1 && !std::getenv("PSASCAN_MAX_BLOCK")) {
    if (max_block >= n && n < (8L << 30)) n_gpus = 1;
    else max_block = std::min(max_block, ((n + n_gpus - 1) / n_gpus + 31) / 32 * 32);
  }
  const long n_blocks = (n + max_block - 1) / max_block;
  n_gpus = (int)std::min<long>(n_gpus, n_blocks);
  gpu.resize(n_gpus);
  // two pinned staging buffers, each at most ram_use / 4 and at most 256 MiB
  size_t stage = (size_t)std::max(1L << 16, std::min(256L << 20, ram_use / 4));

  std::fprintf(stderr, "Input filename = %s\n", input_filename.c_str());
  std::fprintf(stderr, "Output filename = %s\n", output_filename.c_str());
  std::fprintf(stderr, "Gap filename = %s\n", gap_filename.c_str());
  std::fprintf(stderr, "Input length = %ld (%.1LfMiB)\n\n", n, 1.L * n / (1 << 20));
  std::fprintf(stderr, "GPU = %d x %s, HBM free = %.1fGiB\n", n_gpus, gpu_name, free_b / 1073741824.0);
  std::fprintf(stderr, "Host staging = 2 x %.1fMiB (RAM budget %.1LfMiB), threads = %ld (unused)\n",
               stage / 1048576.0, 1.L * ram_use / (1 << 20), max_threads);
  std::fprintf(stderr, "Max block size = %ld (%.1LfMiB), blocks = %ld%s\n\n", max_block,
               1.L * max_block / (1 << 20), n_blocks, em_mode ? " (external-memory mode: SA parts and gap files on disk)" : "");

  for (int g = 1; g < n_gpus; ++g)
    if (psc_init(g, 0, &gpu[g].ctx) != PSC_OK) die_psc("cannot initialise a GPU");
  const long cap = n_gpus > 1 ? psc_dist_slice_capacity(n, n_gpus) : n;
  void *pin = NULL;
  for (int g = 0; g < n_gpus; ++g) {
    if (psc_device_alloc(gpu[g].ctx, (size_t)n, &gpu[g].text_d) != PSC_OK) die_psc("device allocation (text)");
    if (!em_mode && psc_device_alloc(gpu[g].ctx, 5 * (size_t)cap, &gpu[g].out_d) != PSC_OK) die_psc("device allocation (suffix array)");
  }
  if (psc_pinned_alloc(gpu[0].ctx, stage, &pin) != PSC_OK) die_psc("pinned staging buffer");

  // text: file -> pinned buffer -> HBM of every GPU
  std::FILE *f = std::fopen(input_filename.c_str(), "r");
  if (!f) { std::perror(input_filename.c_str()); std::exit(EXIT_FAILURE); }
  for (long done = 0; done < n;) {
    size_t want = (size_t)std::min<long>((long)stage, n - done);
    size_t got = std::fread(pin, 1, want, f);
    if (got != want) {
      std::fprintf(stderr, "\nError: fread in line %s of %s returned %ld\n", "read_text", __FILE__, (long)got);
      std::exit(EXIT_FAILURE);
    }
    for (int g = 0; g < n_gpus; ++g)
      if (psc_copy_h2d(gpu[g].ctx, (char *)gpu[g].text_d + done, pin, got) != PSC_OK) die_psc("host to device copy");
    done += (long)got;
  }
  std::fclose(f);

  const double t_sa = wclock();
  if (em_mode) {
    // partial suffix arrays next to OUTFILE, gap files next to the -g name (a second disk, src/main.cpp:207), merged
    // by streaming: psascan_b200/csrc/em.cu.  SA parts of at most ram_use / 20 bytes as in partial_sufsort.hpp:206.
    const long part_bytes = std::max(4L, ram_use / 20L);
    int rc = psc_construct_sa_files_d(gpu[0].ctx, (const uint8_t *)gpu[0].text_d, n, max_block, output_filename.c_str(),
                                      gap_filename.c_str(), part_bytes);
    if (rc == PSC_ESENTINEL) {
      std::fprintf(stdout, "\n\nError: byte with value 255 was detected in the input text!\n"
                           "See the section on limitations in the README for more information.\n");
      std::fflush(stdout);
      std::exit(EXIT_FAILURE);
    }
    if (rc != PSC_OK) die_psc("suffix array construction (external-memory mode)");
    const double sa_time = wclock() - t_sa, total_time = wclock() - start;
    psc_pinned_free(gpu[0].ctx, pin);
    psc_device_free(gpu[0].ctx, gpu[0].text_d);
    psc_shutdown(gpu[0].ctx);
    std::fprintf(stderr, "\n\nComputation finished. Summary:\n");
    std::fprintf(stderr, "  elapsed time: %.2Lfs (%.4Lfs/MiB)\n", (long double)total_time,
                 (long double)total_time / ((1.L * n) / (1L << 20)));
    std::fprintf(stderr, "  speed: %.2LfMiB/s\n", ((1.L * n) / (1L << 20)) / (long double)total_time);
    std::fprintf(stderr, "  GPU suffix array construction through files: %.3fs (%.1f MB/s)\n", sa_time, n / 1e6 / sa_time);
    return;
  }
  if (n_gpus == 1) {
    gpu[0].rc = psc_construct_sa_d(gpu[0].ctx, (const uint8_t *)gpu[0].text_d, n, max_block, (uint8_t *)gpu[0].out_d);
    gpu[0].err = psc_last_error();
    gpu[0].slice_beg = 0;
    gpu[0].slice_len = n;
  } else {
    uint8_t id[128];
    if (psc_dist_unique_id(id) != PSC_OK) die_psc("NCCL");
    std::vector<std::thread> th;
    for (int g = 0; g < n_gpus; ++g)
      th.emplace_back([&, g] {
        GpuRank &r = gpu[g];
        r.rc = psc_dist_init(r.ctx, id, g, n_gpus);
        if (r.rc == PSC_OK)
          r.rc = psc_construct_sa_dist_slice_d(r.ctx, (const uint8_t *)r.text_d, n, max_block, (uint8_t *)r.out_d,
                                               &r.slice_beg, &r.slice_len);
        if (r.rc != PSC_OK) {
          r.err = psc_last_error();  // the message is per thread
          for (int h = 0; h < n_gpus; ++h)
            if (h != g) psc_dist_abort(gpu[h].ctx);  // the others must not wait for this rank
        }
      });
    for (auto &t : th) t.join();
  }
  for (int g = 0; g < n_gpus; ++g) {
    if (gpu[g].rc == PSC_ESENTINEL) {  // initial_partial_sufsort.hpp:141-146 (message goes to stdout there too)
      std::fprintf(stdout, "\n\nError: byte with value 255 was detected in the input text!\n"
                           "See the section on limitations in the README for more information.\n");
      std::fflush(stdout);
      std::exit(EXIT_FAILURE);
    }
  }
  for (int g = 0; g < n_gpus; ++g)
    if (gpu[g].rc != PSC_OK) {
      std::fprintf(stderr, "\nError: suffix array construction (GPU %d): %s\n", g, gpu[g].err.c_str());
      std::exit(EXIT_FAILURE);
    }
  const double sa_time = wclock() - t_sa;

  // suffix array: HBM -> pinned buffer -> file, slice after slice (slices are in rank order)
  std::FILE *out = std::fopen(output_filename.c_str(), "w");
  if (!out) { std::perror(output_filename.c_str()); std::exit(EXIT_FAILURE); }
  long expect = 0;
  for (int g = 0; g < n_gpus; ++g) {
    if (gpu[g].slice_beg != expect) { std::fprintf(stderr, "\nError: output slices do not tile the suffix array\n"); std::exit(EXIT_FAILURE); }
    expect += gpu[g].slice_len;
    const size_t total = 5 * (size_t)gpu[g].slice_len;
    for (size_t done = 0; done < total;) {
      size_t want = std::min(stage, total - done);
      if (psc_copy_d2h(gpu[g].ctx, pin, (const char *)gpu[g].out_d + done, want) != PSC_OK) die_psc("device to host copy");
      if (std::fwrite(pin, 1, want, out) != want) {
        std::fprintf(stderr, "\nError: fwrite to %s failed\n", output_filename.c_str());
        std::exit(EXIT_FAILURE);
      }
      done += want;
    }
  }
  std::fclose(out);
  if (expect != n) { std::fprintf(stderr, "\nError: output slices do not cover the suffix array\n"); std::exit(EXIT_FAILURE); }

  if (verbose) {
    psc_stats st;
    psc_get_stats(gpu[0].ctx, &st);
    std::fprintf(stderr,
                 "GPU phases of rank 0 (ms): total %.2f | gt_end %.2f | sort %.2f (radix %.2f, %ld passes, %ld rounds) | "
                 "bwt+gt %.2f | rank build %.2f | initial ranks %.2f | LF walk %.2f (%ld steps, %ld walkers) | "
                 "merge %.2f | kernel launches %ld\n",
                 st.ms_total, st.ms_gt_end, st.ms_sort, st.ms_radix, st.radix_passes, st.doubling_rounds, st.ms_bwt,
                 st.ms_rank_build, st.ms_init_ranks, st.ms_walk, st.lf_steps, st.walkers, st.ms_merge,
                 st.kernel_launches);
  }
  psc_pinned_free(gpu[0].ctx, pin);
  for (int g = 0; g < n_gpus; ++g) {
    psc_device_free(gpu[g].ctx, gpu[g].text_d);
    psc_device_free(gpu[g].ctx, gpu[g].out_d);
    psc_dist_finalize(gpu[g].ctx);
    psc_shutdown(gpu[g].ctx);
  }

  // summary in the reference's format (psascan.hpp:128-130)
  const double total_time = wclock() - start;
  std::fprintf(stderr, "\n\nComputation finished. Summary:\n");
  std::fprintf(stderr, "  elapsed time: %.2Lfs (%.4Lfs/MiB)\n", (long double)total_time,
               (long double)total_time / ((1.L * n) / (1L << 20)));
  std::fprintf(stderr, "  speed: %.2LfMiB/s\n", ((1.L * n) / (1L << 20)) / (long double)total_time);
  std::fprintf(stderr, "  GPU suffix array construction: %.3fs (%.1f MB/s) on %d GPU%s\n", sa_time, n / 1e6 / sa_time, n_gpus,
               n_gpus > 1 ? "s" : "");
}

}  // namespace psascan_b200

// The reference's global entry point (psascan.hpp:137-141).
inline void pSAscan(std::string input_filename, std::string output_filename, std::string gap_filename,
                    long ram_use, long max_threads, bool verbose) {
  psascan_b200::pSAscan_impl(input_filename, output_filename, gap_filename, ram_use, max_threads, verbose);
}

#endif
