// host/construct_sa.cpp -- the construct_sa command line of the reference (src/main.cpp:48-246) on
// top of host/psascan.hpp: same options, defaults, messages and exit codes.
//   construct_sa [-g GAPFILE] [-h] [-m MEM] [-o OUTFILE] [-v] FILE
#include <getopt.h>
#include <unistd.h>
#include <cctype>
#include <cstring>
#include <ctime>
#include <string>
#include <thread>

#include "psascan.hpp"

static const char *g_prog = "construct_sa";

static void usage(int status) {
  std::printf(
      "Usage: %s [OPTION]... FILE\n"
      "Construct the suffix array of text stored in FILE.\n"
      "\n"
      "Mandatory arguments to long options are mandatory for short options too.\n"
      "  -h, --help              display this help and exit\n"
      "  -g, --gap=GAPFILE       specify the file holding the gap array. Default:\n"
      "                          OUTFILE.gap, see the -o flag.\n"
      "  -m, --mem=MEM           use MEM bytes of RAM for computation. Metric and IEC\n"
      "                          suffixes are recognized, e.g., -l 10k, -l 1Mi, -l 3G\n"
      "                          gives MEM = 10^4, 2^20, 3*10^6. Default: 3584Mi\n"
      "  -o, --output=OUTFILE    specify output filename. Default: FILE.sa5\n"
      "  -v, --verbose           print detailed information during internal sufsort\n",
      g_prog);
  std::exit(status);
}

// digits followed by an optional case-insensitive K/M/G/T (powers of 1000) or Ki/Mi/Gi/Ti (powers
// of 1024) suffix; anything else is rejected (src/main.cpp:76-131)
static bool parse_size(const char *arg, std::uint64_t *out) {
  std::string s(arg);
  size_t d = 0;
  std::uint64_t v = 0;
  while (d < s.size() && std::isdigit((unsigned char)s[d])) v = v * 10 + (std::uint64_t)(s[d++] - '0');
  if (d == 0) return false;
  std::string suf = s.substr(d);
  for (char &c : suf) c = (char)std::tolower((unsigned char)c);
  if (!suf.empty()) {
    if (suf.size() > 2 || (suf.size() == 2 && suf[1] != 'i')) return false;
    const bool iec = suf.size() == 2;
    int power;
    switch (suf[0]) {
      case 'k': power = 1; break;
      case 'm': power = 2; break;
      case 'g': power = 3; break;
      case 't': power = 4; break;
      default: return false;
    }
    for (int i = 0; i < power; ++i) v = iec ? (v << 10) : (v * 1000);
  }
  *out = v;
  return true;
}

static bool readable(const std::string &name) {
  std::FILE *f = std::fopen(name.c_str(), "r");
  if (f) std::fclose(f);
  return f != NULL;
}

int main(int argc, char **argv) {
  std::srand((unsigned)(std::time(0) + getpid()));
  g_prog = argv[0];
  static const option longopts[] = {{"help", no_argument, NULL, 'h'},       {"gap", required_argument, NULL, 'g'},
                                    {"mem", required_argument, NULL, 'm'},  {"output", required_argument, NULL, 'o'},
                                    {"verbose", no_argument, NULL, 'v'},    {NULL, 0, NULL, 0}};
  std::uint64_t ram_use = (std::uint64_t)3584 << 20;
  std::string out_name, gap_name;
  bool verbose = false;
  for (int c; (c = getopt_long(argc, argv, "g:hm:o:v", longopts, NULL)) != -1;) {
    if (c == 'g') gap_name = optarg;
    else if (c == 'o') out_name = optarg;
    else if (c == 'v') verbose = true;
    else if (c == 'm') {
      if (!parse_size(optarg, &ram_use)) {
        std::fprintf(stderr, "Error: parsing RAM limit (%s) failed\n\n", optarg);
        usage(EXIT_FAILURE);
      }
      if (ram_use == 0) {
        std::fprintf(stderr, "Error: invalid RAM limit (%lu)\n\n", (unsigned long)ram_use);
        usage(EXIT_FAILURE);
      }
    } else usage(EXIT_FAILURE);  // -h included: the reference exits with EXIT_FAILURE there too (:160)
  }
  if (optind >= argc) {
    std::fprintf(stderr, "Error: FILE not provided\n\n");
    usage(EXIT_FAILURE);
  }
  std::string text_name = argv[optind++];
  if (optind < argc)
    std::fprintf(stderr, "Warning: multiple input files provided. Only the first will be processed.\n");
  if (out_name.empty()) out_name = text_name + ".sa5";
  if (gap_name.empty()) gap_name = out_name;  // the code's default (:207); the help text says OUTFILE.gap
  if (!readable(text_name)) {
    std::fprintf(stderr, "Error: input file (%s) does not exist\n\n", text_name.c_str());
    usage(EXIT_FAILURE);
  }
  if (readable(out_name)) {
    char *line = NULL;
    size_t cap = 0;
    ssize_t len;
    do {
      std::printf("Output file (%s) exists. Overwrite? [y/n]: ", out_name.c_str());
      if ((len = getline(&line, &cap, stdin)) == -1) {
        std::printf("\nError: failed to read answer\n\n");
        std::fflush(stdout);
        usage(EXIT_FAILURE);
      }
    } while (len != 2 || (line[0] != 'y' && line[0] != 'n'));
    const bool no = line[0] == 'n';
    std::free(line);
    if (no) std::exit(EXIT_FAILURE);
  }
  long threads = (long)std::thread::hardware_concurrency();  // the reference asks OpenMP (:241)
  if (const char *e = std::getenv("OMP_NUM_THREADS")) threads = std::max(1L, std::atol(e));
  pSAscan(text_name, out_name, gap_name, (long)ram_use, threads, verbose);
  return 0;
}
