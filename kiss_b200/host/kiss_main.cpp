// kiss_main.cpp — the `kISS suffix_sort` command on the B200 path.
//
// Mirrors the reference CLI for this one sub-command: src/main.cpp:19-40 (dispatch), include/utils/options.hpp
// (flags: -h/--help, -v/--version, -g/--generic, -t/--num_threads NUM, --verbose; suffix_sort: -k/--kordered NUM
// default 256, `-k -1` = unbounded, -s/--sorting-algorithm PARALLEL_SORTING|PREFIX_DOUBLING case-insensitive,
// one positional file), include/command/suffix_sort.hpp:21-62 (the command) and include/utils/io.hpp:6-18
// (FASTA when the first byte is '>', else every line is sequence).  Boost.program_options is not available in
// this image, so the parser is hand-written.  `fmindex_build` (include/command/fmindex_build.hpp:14-36) is built too: it
// writes <file>.fmi, byte-compatible with the reference's FMIndex::save, without moving the suffix array off the device;
// `fmindex_query` stays with the reference (it loads that file).
// Extra opt-in flags (defaults unchanged): --gpu IDX, --out FILE (raw little-endian uint32 suffix array —
// the reference never writes it, suffix_sort.hpp:57-61).
#include <algorithm>
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <ctime>
#include <fstream>
#include <iostream>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

#include <biovoltron/algo/sort/kiss1_sorter.hpp>
#include <biovoltron/algo/sort/kiss2_sorter.hpp>

namespace {

const char* kVersion = "1.0.0";  // include/utils/constant.hpp

void usage(std::ostream& os) {
  os << "kISS [--option ...] <command> [--command-option ...] <FASTA filename/Text filename>\n\n"
        "Generic options:\n"
        "  -h [ --help ]             produce help message\n"
        "  -v [ --version ]          print version string\n"
        "  -g [ --generic ]          (Under construction) input contains bases other than ATCG\n"
        "  -t [ --num_threads ] NUM  number of thread (accepted, unused: the GPU path has no host threads)\n"
        "  --verbose                 print more information\n"
        "  --gpu IDX                 CUDA device (default 0)\n"
        "  --gpus N                  shard the LMS sort over N GPUs starting at --gpu (default 1; implies --host-ingest)\n"
        "  --host-ingest             parse the file on the host (the reference's serial getline loop) instead of on the device\n"
        "  --general                 (extension) sort the sequence lines' bytes as they are, any alphabet (proteins, text)\n\n"
        "./kISS suffix_sort [--option ...] <FASTA filename/Text filename>\n\n"
        "Options:\n"
        "  -k [ --kordered ] NUM (=256)          a k-ordered value, where each suffix is sorted based on the\n"
        "                                        first k characters. Using -1 indicates unbounded sorting.\n"
        "  -s [ --sorting-algorithm ] ALGO (=PARALLEL_SORTING)\n"
        "                                        PARALLEL_SORTING or PREFIX_DOUBLING\n"
        "  --out FILE                            write the suffix array (uint32 little endian)\n\n"
        "./kISS fmindex_build <FASTA filename/Text filename>\n"
        "                                        writes <filename>.fmi (FM-index, SA sampling interval 4, k = 32)\n";
}

std::string stamp() {
  using namespace std::chrono;
  auto now = system_clock::now();
  std::time_t t = system_clock::to_time_t(now);
  auto ms = duration_cast<milliseconds>(now.time_since_epoch()).count() % 1000;
  char buf[64], out[96];
  std::strftime(buf, sizeof buf, "%Y-%m-%d %H:%M:%S", std::localtime(&t));
  std::snprintf(out, sizeof out, "%s.%03d", buf, (int)ms);
  return out;
}
// same line shape as the reference's spdlog output (README.md:88, 95-100)
void log_line(const char* level, const char* where, const std::string& msg) {
  std::fprintf(stderr, "[%s] [stderr] [%s] [%s] %s\n", stamp().c_str(), level, where, msg.c_str());
}

// Codec::to_int (istring.hpp:27-46) followed by `% 4` (suffix_sort.hpp:33)
inline std::uint8_t code_of(unsigned char c) {
  switch (c) {
    case 'a': case 'A': return 0;
    case 'c': case 'C': return 1;
    case 'g': case 'G': return 2;
    case 't': case 'T': return 3;
    default: return 0;  // 4 % 4
  }
}

// read_sequence (include/utils/io.hpp:6-18): FASTA mode iff the first byte is '>' — header lines dropped, all
// records concatenated without separators; text mode: every line is sequence.  '\r', N and IUPAC codes all
// become A, as in the reference.  The line after a header is sequence whatever it starts with
// (include/biovoltron/file_io/fasta.hpp:129-133), so a '>' line is a header only when the line before it was not one.
// `raw`: the bytes of the sequence lines as they are (--general), instead of their base codes
std::vector<std::uint8_t> read_sequence(std::istream& fin, bool raw = false) {
  std::vector<std::uint8_t> seq;
  bool fasta = fin.peek() == '>', prev_header = false;
  std::string line;
  while (std::getline(fin, line)) {
    const bool header = fasta && !line.empty() && line[0] == '>' && !prev_header;
    prev_header = header;
    if (header) continue;
    for (unsigned char c : line) seq.push_back(raw ? c : code_of(c));
  }
  return seq;
}

struct Args {
  bool help = false, version = false, generic = false, verbose = false, host_ingest = false, general = false;
  std::size_t threads = std::thread::hardware_concurrency();
  int gpu = 0, gpus = 1;
  std::string command, file, out;
  std::size_t k = 256;
  int algo = KISS_B200_PARALLEL_SORTING;
};

std::string upper(std::string s) {
  std::transform(s.begin(), s.end(), s.begin(), [](unsigned char c) { return (char)std::toupper(c); });
  return s;
}

// "-1" lexically converts to 2^64-1 for a size_t option; the call narrows it to 0xFFFFFFFF
std::size_t parse_size(const std::string& v) {
  if (v.empty()) throw std::invalid_argument("empty number");
  std::size_t idx = 0;
  if (v[0] == '-') {
    long long s = std::stoll(v, &idx);
    if (idx != v.size()) throw std::invalid_argument("bad number " + v);
    return (std::size_t)s;
  }
  unsigned long long u = std::stoull(v, &idx);
  if (idx != v.size()) throw std::invalid_argument("bad number " + v);
  return (std::size_t)u;
}

Args parse(int argc, char** argv) {
  Args a;
  std::vector<std::string> pos;
  auto need = [&](int& i, const std::string& flag) -> std::string {
    if (i + 1 >= argc) throw std::invalid_argument("the required argument for option '" + flag + "' is missing");
    return argv[++i];
  };
  for (int i = 1; i < argc; i++) {
    std::string s = argv[i];
    auto eq = s.find('=');
    std::string flag = (s.rfind("--", 0) == 0 && eq != std::string::npos) ? s.substr(0, eq) : s;
    auto value = [&](const std::string& f) { return flag != s ? s.substr(eq + 1) : need(i, f); };
    if (flag == "-h" || flag == "--help") a.help = true;
    else if (flag == "-v" || flag == "--version") a.version = true;
    else if (flag == "-g" || flag == "--generic") a.generic = true;
    else if (flag == "--general") a.general = true;
    else if (flag == "--verbose") a.verbose = true;
    else if (flag == "-t" || flag == "--num_threads") a.threads = parse_size(value(flag));
    else if (flag == "--gpu") a.gpu = (int)parse_size(value(flag));
    else if (flag == "--gpus") a.gpus = (int)parse_size(value(flag));
    else if (flag == "--out") a.out = value(flag);
    else if (flag == "--host-ingest") a.host_ingest = true;
    else if (flag == "-k" || flag == "--kordered") a.k = parse_size(value(flag));
    else if (flag == "-s" || flag == "--sorting-algorithm") {
      std::string v = upper(value(flag));  // options.hpp:67-81
      if (v == "PARALLEL_SORTING") a.algo = KISS_B200_PARALLEL_SORTING;
      else if (v == "PREFIX_DOUBLING") a.algo = KISS_B200_PREFIX_DOUBLING;
      else throw std::invalid_argument("the argument for option '--sorting-algorithm' is invalid");
    } else if (s.size() > 1 && s[0] == '-' && !(s.size() > 1 && std::isdigit((unsigned char)s[1])))
      throw std::invalid_argument("unrecognised option '" + s + "'");
    else pos.push_back(s);
  }
  if (!pos.empty()) a.command = pos[0];
  if (pos.size() > 1) a.file = pos[1];
  if (pos.size() > 2) throw std::invalid_argument("too many positional options have been specified on the command line");
  return a;
}

int suffix_sort_main(const Args& a) {
  if (a.generic)  // suffix_sort.hpp:26-28
    throw std::invalid_argument("Generic sorting and indexing are currently not supported.");
  if (a.file.empty()) throw std::invalid_argument("the option '--fasta' is required but missing");
  std::ifstream fa(a.file, std::ios::binary);
  if (!fa) throw std::invalid_argument("cannot open " + a.file);
  biovoltron::kiss::b200::device_index() = a.gpu;
  if (a.gpus > 1) biovoltron::kiss::b200::gpus() = a.gpus;
  auto k32 = (std::uint32_t)a.k;  // narrows at the call like suffix_sort.hpp:59
  biovoltron::kiss::vector<std::uint32_t> sa;
  std::uint64_t n_text = 0;
  double sec = 0;
  if (a.general) {
    // extension (not in the reference's CLI, whose -g is refused above): the sequence lines' bytes as they are — a protein
    // FASTA, any text — through the general-alphabet sorters (kiss1_sorter.hpp:28-44 -> kiss_b200_suffix_sort_u32)
    auto seq = read_sequence(fa, true);
    n_text = seq.size();
    auto t0 = std::chrono::steady_clock::now();
    if (a.algo == KISS_B200_PARALLEL_SORTING) sa = biovoltron::KISS1Sorter<std::uint32_t>::get_suffix_array(seq, k32, a.threads);
    else sa = biovoltron::KISS2Sorter<std::uint32_t>::get_suffix_array(seq, k32, a.threads);
    sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  } else if (a.host_ingest || a.gpus > 1) {
    // the reference's way: parse on the host (serial getline), then the sorter facade on the codes (the facade shards the
    // LMS sort over --gpus N GPUs: kiss_b200_suffix_sort_dna_u32_multi)
    auto seq = read_sequence(fa);
    n_text = seq.size();
    auto t0 = std::chrono::steady_clock::now();
    if (a.algo == KISS_B200_PARALLEL_SORTING) sa = biovoltron::KISS1Sorter<std::uint32_t>::get_suffix_array_dna(seq, k32, a.threads);
    else sa = biovoltron::KISS2Sorter<std::uint32_t>::get_suffix_array_dna(seq, k32, a.threads);
    sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  } else {
    // default: the file's bytes go to the device as they are; header / newline stripping and base encoding run
    // there (read_sequence semantics, ingest.cu) — no per-character host loop
    fa.seekg(0, std::ios::end);
    const std::streamoff bytes = fa.tellg();
    fa.seekg(0, std::ios::beg);
    std::vector<std::uint8_t> file((std::size_t)(bytes > 0 ? bytes : 0));
    if (!file.empty()) fa.read(reinterpret_cast<char*>(file.data()), (std::streamsize)file.size());
    if (!fa && !file.empty()) throw std::runtime_error("cannot read " + a.file);
    auto t0 = std::chrono::steady_clock::now();
    n_text = biovoltron::kiss::b200::suffix_array_file(file, sa, k32, a.algo);
    sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  }
  if (a.verbose && n_text) {  // the reference's per-stage DEBUG lines (kiss1_core.hpp:244-267; README.md:95-100)
    const auto& t = biovoltron::kiss::b200::last_timings();
    const char* w = "kiss_b200";
    log_line("debug", w, "host to device elapsed " + std::to_string(t.h2d));
    log_line("debug", w, "pack elapsed " + std::to_string(t.pack));
    log_line("debug", w, "get_lms elapsed " + std::to_string(t.get_lms));
    log_line("debug", w, "lms_suffix_direct_sort elapsed " + std::to_string(t.lms_sort));
    log_line("debug", w, "put_lms_suffix elapsed " + std::to_string(t.put_lms));
    log_line("debug", w, "induced_sort elapsed " + std::to_string(t.induce));
    log_line("debug", w, "device to host elapsed " + std::to_string(t.d2h));
  }
  log_line("info", "suffix_sort.hpp:61",
           "n = " + std::to_string(n_text) + ", k = " + std::to_string(a.k) + ", suffix sorting elapsed " + std::to_string(sec));
  if (!a.out.empty()) {
    std::ofstream os(a.out, std::ios::binary);
    os.write(reinterpret_cast<const char*>(sa.data()), (std::streamsize)(sa.size() * sizeof(std::uint32_t)));
    if (!os) throw std::runtime_error("cannot write " + a.out);
  }
  return 0;
}

// fmindex_build_main (include/command/fmindex_build.hpp:14-36): read_sequence, `% 4`, FMIndex<4, uint32_t,
// KISS1Sorter<uint32_t>>{.LOOKUP_LEN = 0}.build (k = 32, fm_index.hpp:385-386), save to <fasta>.fmi
int fmindex_build_main(const Args& a) {
  if (a.generic) throw std::invalid_argument("Generic sorting and indexing are currently not supported.");
  if (a.file.empty()) throw std::invalid_argument("the option '--fasta' is required but missing");
  std::ifstream fa(a.file, std::ios::binary);
  if (!fa) throw std::invalid_argument("cannot open " + a.file);
  biovoltron::kiss::b200::device_index() = a.gpu;
  fa.seekg(0, std::ios::end);
  const std::streamoff bytes = fa.tellg();
  fa.seekg(0, std::ios::beg);
  std::vector<std::uint8_t> file((std::size_t)(bytes > 0 ? bytes : 0)), image;
  if (!file.empty()) fa.read(reinterpret_cast<char*>(file.data()), (std::streamsize)file.size());
  if (!fa && !file.empty()) throw std::runtime_error("cannot read " + a.file);
  double fm = 0;
  auto t0 = std::chrono::steady_clock::now();
  const std::uint64_t n = biovoltron::kiss::b200::fmindex_image_file(file, image, 32u, KISS_B200_PARALLEL_SORTING, 4u, &fm);
  const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  std::ofstream os(a.file + ".fmi", std::ios::binary);
  os.write(reinterpret_cast<const char*>(image.data()), (std::streamsize)image.size());
  if (!os) throw std::runtime_error("cannot write " + a.file + ".fmi");
  if (a.verbose) {
    const auto& t = biovoltron::kiss::b200::last_timings();
    log_line("debug", "kiss_b200", "suffix array (device) elapsed " + std::to_string(t.total_device));
    log_line("debug", "kiss_b200", "bwt + occ + sampled sa (device) elapsed " + std::to_string(fm));
  }
  log_line("info", "fmindex_build", "n = " + std::to_string(n) + ", " + std::to_string(image.size()) + " bytes, index building elapsed " + std::to_string(sec));
  return 0;
}

}  // namespace

int main(int argc, char** argv) {
  Args a;
  try {
    a = parse(argc, argv);
  } catch (const std::exception& e) {  // options.hpp:240-244: message, help, exit(1)
    std::cerr << e.what() << "\n";
    usage(std::cerr);
    return 1;
  }
  if (a.version) {  // options.hpp:255-264
    std::cout << kVersion << "\n";
    return 1;
  }
  if (a.help || a.command.empty()) {
    usage(std::cout);
    return 1;
  }
  try {
    if (a.command == "suffix_sort") return suffix_sort_main(a);
    if (a.command == "fmindex_build") return fmindex_build_main(a);
    if (a.command == "fmindex_query")
      throw std::invalid_argument("command 'fmindex_query' is outside the B200 path (the reference's own loads the .fmi written here)");
    throw std::invalid_argument("the argument ('" + a.command + "') for option '--command' is invalid");  // options.hpp:298-299
  } catch (const std::exception& e) {
    std::cerr << "terminate called after throwing: " << e.what() << "\n";
    return 134;  // the reference lets the exception escape main (abort)
  }
}
