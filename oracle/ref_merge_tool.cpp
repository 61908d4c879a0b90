// ref_merge_tool.cpp -- TEST INFRASTRUCTURE ONLY (like everything under oracle/).
// Runs the REFERENCE's own merge() (/root/reference/include/merge.hpp:54-180, compiled from where it lies) on
// partial-suffix-array part files and v-byte gap files that the GPU library wrote in external-memory mode
// (psascan_b200/csrc/em.cu): if the reference's readers (io/distributed_file.hpp, io/async_vbyte_stream_reader.hpp)
// and its merge produce the right suffix array from those files, the formats are the reference's.
//   ref_merge MANIFEST OUT.sa5
// MANIFEST (written by the library when PSC_KEEP_PARTS is set): one line per block, left to right:
//   beg end sa_base items_per_part n_parts gap_file|- gt_begin_file|-   (the last field is not used here)
// The reference's merge deletes every file it has read (merge.hpp:178-179): run it on copies.
#include <cstdio>
#include <fstream>
#include <sstream>
#include <string>
#include <vector>
#include "merge.hpp"

using namespace psascan_private;

int main(int argc, char **argv) {
  if (argc != 3) { std::fprintf(stderr, "usage: %s MANIFEST OUT.sa5\n", argv[0]); return 2; }
  std::ifstream in(argv[1]);
  std::vector<half_block_info<int> > info;
  std::string line;
  while (std::getline(in, line)) {
    if (line.empty()) continue;
    std::istringstream ss(line);
    long beg, end, items_per_part, n_parts;
    std::string sa_base, gap_file;
    ss >> beg >> end >> sa_base >> items_per_part >> n_parts >> gap_file;
    half_block_info<int> h;
    h.beg = beg;
    h.end = end;
    h.gap_filename = gap_file == "-" ? std::string() : gap_file;
    // a distributed_file that is already "written": the object only needs the name, the item counts and the state
    h.psa = new distributed_file<int>("unused", items_per_part * (long)sizeof(int));
    h.psa->m_filename = sa_base;
    h.psa->m_state = distributed_file<int>::STATE_WRITTEN;
    h.psa->m_total_write = end - beg;
    h.psa->m_files_cnt = n_parts;
    info.push_back(h);
  }
  if (info.empty()) { std::fprintf(stderr, "empty manifest\n"); return 2; }
  merge<int>(argv[2], 64L << 20, info);
  return 0;
}
