// proton_em_cfg.cpp — Level-1 entry points: what the reference PROCESS does for one analytical.cfg (proton_em_run_cfg),
// the same for many cfgs in one batched evaluation (proton_em_run_cfgs), and the packing of a whole project directory
// from the files the reference's formulation wrote (proton_em_pack_from_project).
//
// Mirrors main / simulation_engine::read_file / analyical_wrapper / tripletParser of the reference
// binary (bin/EMtool_analytical@0x6920, @0xaaf4, @0xb93c, @0xb010), including
// the texts it prints to stdout before exit(1) (captured by running the binary, tests/test_launcher_cpu.py).
// The arithmetic itself goes through proton_em_pack / proton_em_eval on the GPU.
//
// Deliberate differences (the reference aborts with SIGABRT in these cases; we report an error, rc 1):
//   * a B.csv index outside the matrix (Eigen assertion in set_from_triplets),
//   * a curden.csv line that std::stod cannot parse, or a missing curden.csv,
//   * an input_files path that does not contain "input".
#include "../../include/proton_em.h"

#include <algorithm>
#include <atomic>
#include <cerrno>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <dirent.h>
#include <fstream>
#include <new>
#include <sstream>
#include <string>
#include <sys/stat.h>
#include <thread>
#include <vector>

namespace {

void set_msg(char* msg, size_t n, const std::string& s) {
  if (!msg || n == 0) return;
  std::snprintf(msg, n, "%s", s.c_str());
}

bool make_dirs(const std::string& path) {
  std::string cur;
  for (size_t i = 0; i <= path.size(); ++i) {
    if (i == path.size() || path[i] == '/') {
      if (!cur.empty() && cur != "/") {
        if (mkdir(cur.c_str(), 0777) != 0 && errno != EEXIST) return false;
      }
    }
    if (i < path.size()) cur.push_back(path[i]);
  }
  return true;
}

std::vector<std::string> tokenize(const std::string& line) {   // boost::char_separator<char>(", \t")
  std::vector<std::string> out;
  std::string cur;
  for (char c : line) {
    if (c == ',' || c == ' ' || c == '\t') {
      if (!cur.empty()) { out.push_back(cur); cur.clear(); }
    } else {
      cur.push_back(c);
    }
  }
  if (!cur.empty()) out.push_back(cur);
  return out;
}

// What one analytical.cfg names: the line's inputs as the reference process reads them.
struct CfgLine {
  std::string out_dir;
  int32_t N = 0;
  double a_coeff = 0.0, sim_time = 0.0;
  std::vector<double> u;
  std::vector<int32_t> rows, cols;
  std::vector<double> vals;
};

// The reading half of analyical_wrapper: output directory, curden.csv, B.csv of the directory `input_files` (with its
// trailing '/'); out.N must be set.  On failure `text` is what the reference prints before exit(1).
int read_line_files(const std::string& input_files, CfgLine& out, std::string& text) {
  const long num_nodes = out.N;
  const size_t pos = input_files.find("input");
  if (pos == std::string::npos) { text = "The input directory path does not contain \"input\".\n"; return PROTON_EM_EINVAL; }
  out.out_dir = input_files;
  out.out_dir.replace(pos, 5, "output");
  {
    const std::string path = input_files + "curden.csv";
    std::ifstream f(path);
    if (!f) { text = "The file " + path + " was not found.\nTerminating.\n"; return PROTON_EM_EIO; }
    std::string line;
    size_t lineno = 0;
    while (std::getline(f, line)) {
      ++lineno;
      char* end = nullptr;
      errno = 0;
      const double v = std::strtod(line.c_str(), &end);
      if (end == line.c_str()) {
        text = "Error parsing file " + path + " at line " + std::to_string(lineno) + ".\nTerminating.\n";
        return PROTON_EM_EINVAL;
      }
      out.u.push_back(v);
    }
  }
  {
    const std::string path = input_files + "B.csv";
    std::ifstream f(path);
    if (!f) { text = "Terminating\nThe file " + path + " was not found.\nTerminating.\n"; return PROTON_EM_EIO; }
    std::string line;
    size_t lineno = 0;
    while (std::getline(f, line)) {
      ++lineno;
      const std::vector<std::string> tok = tokenize(line);
      if (tok.empty()) continue;
      if (tok.size() < 2) { text = "Required second value at line " + std::to_string(lineno) + " of the file " + path + ".\nTerminating.\n"; return PROTON_EM_EINVAL; }
      if (tok.size() < 3) { text = "Required third value at line " + std::to_string(lineno) + " of the file " + path + ".\nTerminating.\n"; return PROTON_EM_EINVAL; }
      char* e1 = nullptr; char* e2 = nullptr;
      const long r = std::strtol(tok[0].c_str(), &e1, 10), c = std::strtol(tok[1].c_str(), &e2, 10);
      if (e1 == tok[0].c_str() || e2 == tok[1].c_str()) {
        text = "Error parsing file " + path + " at line " + std::to_string(lineno) + ".\nTerminating.\n";
        return PROTON_EM_EINVAL;
      }
      out.rows.push_back((int32_t)r); out.cols.push_back((int32_t)c); out.vals.push_back(std::atof(tok[2].c_str()));
    }
  }
  for (size_t i = 0; i < out.rows.size(); ++i) {
    if (out.rows[i] < 0 || out.rows[i] >= num_nodes || out.cols[i] < 0 || (size_t)out.cols[i] >= out.u.size()) {
      text = "B.csv entry " + std::to_string(i + 1) + " lies outside the " + std::to_string(num_nodes) + " x " +
             std::to_string(out.u.size()) + " matrix.\nTerminating.\n";
      return PROTON_EM_EINVAL;
    }
  }
  return PROTON_EM_OK;
}


// read_file + the reading half of analyical_wrapper; on failure `text` is what the reference prints before exit(1).
int read_cfg_line(const char* cfg_path, CfgLine& out, std::string& text) {
  if (!cfg_path) { text = "Wrong number of arguments. Configuration file expected.\n"; return PROTON_EM_EINVAL; }
  // ---- read_file: "key value" per line, later keys override, unknown keys ignored -----------------
  std::string input_files;
  long num_nodes = 0;
  double a_coeff = 0.0, sim_time = 0.0;
  bool have_in = false, have_n = false, have_a = false, have_t = false;
  {
    std::ifstream f(cfg_path);
    std::string line;
    while (f && std::getline(f, line)) {
      std::istringstream is(line);
      std::string key;
      if (!(is >> key)) continue;
      if (key == "input_files") { std::string v; if (is >> v) { input_files = v; have_in = true; } }
      else if (key == "num_nodes") { long v; if (is >> v) { num_nodes = v; have_n = true; } }
      else if (key == "A_coeff") { double v; if (is >> v) { a_coeff = v; have_a = true; } }
      else if (key == "simulation_time") { double v; if (is >> v) { sim_time = v; have_t = true; } }
    }
  }
  if (!have_in) { text = std::string(cfg_path) + " The configuration file does not include the input directory.\n"; return PROTON_EM_EINVAL; }
  if (!have_n) { text = "The configuration file does not include the number of nodes.\n"; return PROTON_EM_EINVAL; }
  if (!have_a) { text = "The configuration file does not include the A coefficient.\n"; return PROTON_EM_EINVAL; }
  if (!have_t) { text = "The configuration file does not include the target simulation time.\n"; return PROTON_EM_EINVAL; }
  if (num_nodes <= 0 || num_nodes > 0x7fffffffL) { text = "num_nodes must be a positive 32-bit integer.\n"; return PROTON_EM_EINVAL; }

  out.N = (int32_t)num_nodes; out.a_coeff = a_coeff; out.sim_time = sim_time;
  return read_line_files(input_files, out, text);
}

// stress_<%f>.txt, one "%.18g" per line (the writing half of analyical_wrapper)
int write_stress_file(const CfgLine& ln, const double* stress, std::string& text) {
  const std::string fname = ln.out_dir + "stress_" + std::to_string(ln.sim_time) + ".txt";
  const size_t slash = fname.find_last_of('/');
  if (slash != std::string::npos && !make_dirs(fname.substr(0, slash))) {
    text = "Cannot create the output directory " + fname.substr(0, slash) + ".\nTerminating.\n";
    return PROTON_EM_EIO;
  }
  FILE* fo = std::fopen(fname.c_str(), "w");
  if (!fo) { text = "Cannot open " + fname + " for writing.\nTerminating.\n"; return PROTON_EM_EIO; }
  for (int32_t i = 0; i < ln.N; ++i) std::fprintf(fo, "%.18g\n", stress[(size_t)i]);
  std::fclose(fo);
  return PROTON_EM_OK;
}

}  // namespace

// No exception leaves the C ABI: the entry points below are function-try-blocks that end in this.
#define EM_CFG_CATCH                                                                                                  \
  catch (const std::bad_alloc&) { set_msg(msg, msg_len, "Out of host memory.\n"); return PROTON_EM_ENOMEM; }          \
  catch (const std::exception& e) { set_msg(msg, msg_len, std::string(e.what()) + "\n"); return PROTON_EM_EINVAL; }

extern "C" int proton_em_run_cfg(proton_em_ctx* ctx, const char* cfg_path, char* msg, size_t msg_len) try {
  set_msg(msg, msg_len, "");
  CfgLine ln;
  std::string text;
  { const int rc = read_cfg_line(cfg_path, ln, text); if (rc) { set_msg(msg, msg_len, text); return rc; } }
  // ---- analytical(): one line, one timestamp on the GPU -----------------------------------------
  if (!ctx) { set_msg(msg, msg_len, "No CUDA context.\n"); return PROTON_EM_ECUDA; }
  const int64_t trip_off[2] = {0, (int64_t)ln.rows.size()};
  const int64_t u_off[2] = {0, (int64_t)ln.u.size()};
  proton_em_batch batch{1, &ln.N, &ln.a_coeff, trip_off, ln.rows.data(), ln.cols.data(), ln.vals.data(), u_off, ln.u.data()};
  std::vector<double> stress((size_t)ln.N);
  double mx = 0.0; int32_t amx = 0; uint8_t flag = 0;
  proton_em_result res{stress.data(), &mx, &amx, &flag};
  proton_em_opts opts{0.0, PROTON_EM_ALG_AUTO, 0};
  const int rc = proton_em_analyze_host(ctx, &batch, &ln.sim_time, 1, &opts, &res, 0);
  if (rc != PROTON_EM_OK) { set_msg(msg, msg_len, std::string(proton_em_last_error()) + "\nTerminating.\n"); return rc; }
  const int wrc = write_stress_file(ln, stress.data(), text);
  if (wrc) set_msg(msg, msg_len, text);
  return wrc;
} EM_CFG_CATCH

// Many analytical.cfg files in one call — what PROTON_cli.analyze starts as one process per line
// (PROTON_cli.py:1174-1181).  Every cfg is read like proton_em_run_cfg reads it; the lines that read cleanly form ONE
// batch, evaluated once at the distinct simulation times they ask for, and every line's stress_<t>.txt is written where
// the reference process would write it.  status[i] receives the code of cfg i (0 ok); the return value is the number of
// cfgs that failed, or a negative PROTON_EM_E* code if the batched evaluation itself failed (message in `msg`, which
// otherwise holds the reference's stdout text of the first failing cfg).
extern "C" int proton_em_run_cfgs(proton_em_ctx* ctx, const char* const* cfg_paths, int64_t n, int32_t* status,
                                  char* msg, size_t msg_len) try {
  set_msg(msg, msg_len, "");
  if (n < 0 || (n > 0 && !cfg_paths)) { set_msg(msg, msg_len, "proton_em_run_cfgs: bad arguments.\n"); return PROTON_EM_EINVAL; }
  if (!ctx) { set_msg(msg, msg_len, "No CUDA context.\n"); return PROTON_EM_ECUDA; }
  std::vector<CfgLine> lines((size_t)n);
  std::vector<int64_t> good;
  int failed = 0;
  bool have_msg = false;
  for (int64_t i = 0; i < n; ++i) {
    std::string text;
    const int rc = read_cfg_line(cfg_paths[i], lines[(size_t)i], text);
    if (status) status[i] = rc;
    if (rc) { ++failed; if (!have_msg) { set_msg(msg, msg_len, text); have_msg = true; } }
    else good.push_back(i);
  }
  if (good.empty()) return failed;
  // one batch over the good lines, evaluated at the distinct times
  std::vector<double> ts;
  std::vector<int> t_of(good.size());
  for (size_t g = 0; g < good.size(); ++g) {
    const double t = lines[(size_t)good[g]].sim_time;
    size_t k = 0;
    while (k < ts.size() && ts[k] != t) ++k;
    if (k == ts.size()) ts.push_back(t);
    t_of[g] = (int)k;
  }
  const int64_t L = (int64_t)good.size();
  const int32_t T = (int32_t)ts.size();
  std::vector<int32_t> nn((size_t)L), rows, cols;
  std::vector<double> aa((size_t)L), vals, u;
  std::vector<int64_t> trip_off((size_t)L + 1, 0), u_off((size_t)L + 1, 0), node_off((size_t)L + 1, 0);
  for (int64_t g = 0; g < L; ++g) {
    const CfgLine& ln = lines[(size_t)good[(size_t)g]];
    nn[(size_t)g] = ln.N; aa[(size_t)g] = ln.a_coeff;
    rows.insert(rows.end(), ln.rows.begin(), ln.rows.end());
    cols.insert(cols.end(), ln.cols.begin(), ln.cols.end());
    vals.insert(vals.end(), ln.vals.begin(), ln.vals.end());
    u.insert(u.end(), ln.u.begin(), ln.u.end());
    trip_off[(size_t)g + 1] = (int64_t)rows.size(); u_off[(size_t)g + 1] = (int64_t)u.size();
    node_off[(size_t)g + 1] = node_off[(size_t)g] + ln.N;
  }
  proton_em_batch batch{L, nn.data(), aa.data(), trip_off.data(), rows.data(), cols.data(), vals.data(), u_off.data(), u.data()};
  std::vector<double> stress((size_t)node_off[(size_t)L] * (size_t)T), mx((size_t)L * T);
  std::vector<int32_t> amx((size_t)L * T);
  std::vector<uint8_t> flag((size_t)L * T);
  proton_em_result res{stress.data(), mx.data(), amx.data(), flag.data()};
  proton_em_opts opts{0.0, PROTON_EM_ALG_AUTO, 0};
  const int rc = proton_em_analyze_host(ctx, &batch, ts.data(), T, &opts, &res, 0);
  if (rc != PROTON_EM_OK) { set_msg(msg, msg_len, std::string(proton_em_last_error()) + "\nTerminating.\n"); return rc; }
  for (int64_t g = 0; g < L; ++g) {
    const CfgLine& ln = lines[(size_t)good[(size_t)g]];
    std::string text;
    const double* row = stress.data() + (size_t)node_off[(size_t)g] * (size_t)T + (size_t)t_of[(size_t)g] * (size_t)ln.N;
    const int wrc = write_stress_file(ln, row, text);
    if (wrc) {
      if (status) status[good[(size_t)g]] = wrc;
      ++failed;
      if (!have_msg) { set_msg(msg, msg_len, text); have_msg = true; }
    }
  }
  return failed;
} EM_CFG_CATCH

// ---------------------------------------------------------------------------------------------------------------------
// proton_em_pack_from_project: every discretised line of a PROTON project directory in one plan.
// Matrix_Formulation_Class leaves, per line and per (technology, temperature, width) variant, the directory
//   <project>/input/<line>/<technology>_<temperature>_<width>/{analytical.txt, B.csv, curden.csv}
// (matrix_formulation_func_cli.py:217-286); Analytical_Class reads nx_total and Acoef from analytical.txt
// (analytical_class_cli.py:31-41: `words = line.split(); if 'Acoef' in words: Acoef = words[2]`) and the process reads
// the two csv files.  The files are parsed on a few host threads; the arithmetic starts with proton_em_pack.
// ---------------------------------------------------------------------------------------------------------------------
namespace {

std::vector<std::string> list_dirs(const std::string& path) {
  std::vector<std::string> out;
  DIR* d = opendir(path.c_str());
  if (!d) return out;
  while (dirent* e = readdir(d)) {
    const std::string name = e->d_name;
    if (name == "." || name == "..") continue;
    struct stat st;
    if (stat((path + "/" + name).c_str(), &st) == 0 && S_ISDIR(st.st_mode)) out.push_back(name);
  }
  closedir(d);
  std::sort(out.begin(), out.end());
  return out;
}

bool file_exists(const std::string& path) {
  struct stat st;
  return stat(path.c_str(), &st) == 0 && S_ISREG(st.st_mode);
}

// analytical.txt: "nx_total = <int>" and "Acoef = <float>", any order, whitespace-separated words
int read_analytical_txt(const std::string& path, CfgLine& out) {
  std::ifstream f(path);
  if (!f) return PROTON_EM_EIO;
  bool have_n = false, have_a = false;
  std::string line;
  while (std::getline(f, line)) {
    std::istringstream is(line);
    std::vector<std::string> words;
    for (std::string w; is >> w;) words.push_back(w);
    if (words.size() < 3) continue;
    for (const std::string& w : words) {
      if (w == "Acoef") { out.a_coeff = std::atof(words[2].c_str()); have_a = true; }
      if (w == "nx_total") {
        const long v = std::strtol(words[2].c_str(), nullptr, 10);
        if (v <= 0 || v > 0x7fffffffL) return PROTON_EM_EINVAL;
        out.N = (int32_t)v; have_n = true;
      }
    }
  }
  return (have_n && have_a) ? PROTON_EM_OK : PROTON_EM_EINVAL;
}

}  // namespace

extern "C" int proton_em_pack_from_project(proton_em_ctx* ctx, const char* project_dir, const char* variant,
                                           proton_em_plan** out, int64_t* skipped, char* msg, size_t msg_len) try {
  set_msg(msg, msg_len, "");
  if (skipped) *skipped = 0;
  if (!out || !project_dir) { set_msg(msg, msg_len, "proton_em_pack_from_project: bad arguments.\n"); return PROTON_EM_EINVAL; }
  *out = nullptr;
  if (!ctx) { set_msg(msg, msg_len, "No CUDA context.\n"); return PROTON_EM_ECUDA; }
  const std::string in_root = std::string(project_dir) + "/input";
  const std::vector<std::string> line_names = list_dirs(in_root);
  if (line_names.empty()) {
    set_msg(msg, msg_len, "No discretization has been performed in " + in_root + ".\n");
    return PROTON_EM_EIO;
  }
  const std::string want = variant ? variant : "";
  const size_t n = line_names.size();
  std::vector<CfgLine> lines(n);
  std::vector<int> rc(n, PROTON_EM_OK);
  std::vector<std::string> text(n);
  std::atomic<size_t> next{0};
  auto work = [&]() {
    for (size_t i = next.fetch_add(1); i < n; i = next.fetch_add(1)) {
      const std::string line_dir = in_root + "/" + line_names[i];
      std::string var = want;
      if (var.empty()) {
        for (const std::string& v : list_dirs(line_dir))
          if (file_exists(line_dir + "/" + v + "/analytical.txt")) { var = v; break; }
      }
      const std::string dir = line_dir + "/" + var + "/";
      if (var.empty() || read_analytical_txt(dir + "analytical.txt", lines[i]) != PROTON_EM_OK) {
        rc[i] = PROTON_EM_EIO;
        text[i] = "No discretization has been performed for line " + line_names[i] +
                  (want.empty() ? std::string(".\n") : " with params: " + want + ".\n");
        continue;
      }
      rc[i] = read_line_files(dir, lines[i], text[i]);
    }
  };
  {
    unsigned nt = std::thread::hardware_concurrency();
    nt = nt == 0 ? 1 : (nt > 16 ? 16 : nt);
    if (nt > n) nt = (unsigned)n;
    std::vector<std::thread> pool;
    for (unsigned t = 1; t < nt; ++t) pool.emplace_back(work);
    work();
    for (std::thread& t : pool) t.join();
  }
  std::vector<int32_t> nn, rows, cols;
  std::vector<double> aa, vals, u;
  std::vector<int64_t> trip_off(1, 0), u_off(1, 0);
  std::vector<const char*> names;
  int64_t bad = 0;
  for (size_t i = 0; i < n; ++i) {
    if (rc[i] != PROTON_EM_OK) {
      if (bad++ == 0) set_msg(msg, msg_len, text[i]);
      continue;
    }
    const CfgLine& ln = lines[i];
    nn.push_back(ln.N); aa.push_back(ln.a_coeff);
    rows.insert(rows.end(), ln.rows.begin(), ln.rows.end());
    cols.insert(cols.end(), ln.cols.begin(), ln.cols.end());
    vals.insert(vals.end(), ln.vals.begin(), ln.vals.end());
    u.insert(u.end(), ln.u.begin(), ln.u.end());
    trip_off.push_back((int64_t)rows.size()); u_off.push_back((int64_t)u.size());
    names.push_back(line_names[i].c_str());
  }
  if (skipped) *skipped = bad;
  if (nn.empty()) return PROTON_EM_EIO;
  proton_em_batch batch{(int64_t)nn.size(), nn.data(), aa.data(), trip_off.data(), rows.data(), cols.data(), vals.data(),
                        u_off.data(), u.data()};
  const int prc = proton_em_pack(ctx, &batch, PROTON_EM_HOST, out);
  if (prc != PROTON_EM_OK) { set_msg(msg, msg_len, std::string(proton_em_last_error()) + "\n"); return prc; }
  return proton_em_plan_set_line_names(*out, names.data(), (int64_t)names.size());
} EM_CFG_CATCH
