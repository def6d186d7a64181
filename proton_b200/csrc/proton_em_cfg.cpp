// proton_em_cfg.cpp — Level-1 entry point: what the reference PROCESS does for one analytical.cfg.
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

#include <cerrno>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <string>
#include <sys/stat.h>
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

}  // namespace

extern "C" int proton_em_run_cfg(proton_em_ctx* ctx, const char* cfg_path, char* msg, size_t msg_len) {
  set_msg(msg, msg_len, "");
  if (!cfg_path) { set_msg(msg, msg_len, "Wrong number of arguments. Configuration file expected.\n"); return PROTON_EM_EINVAL; }
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
  if (!have_in) { set_msg(msg, msg_len, std::string(cfg_path) + " The configuration file does not include the input directory.\n"); return PROTON_EM_EINVAL; }
  if (!have_n) { set_msg(msg, msg_len, "The configuration file does not include the number of nodes.\n"); return PROTON_EM_EINVAL; }
  if (!have_a) { set_msg(msg, msg_len, "The configuration file does not include the A coefficient.\n"); return PROTON_EM_EINVAL; }
  if (!have_t) { set_msg(msg, msg_len, "The configuration file does not include the target simulation time.\n"); return PROTON_EM_EINVAL; }
  if (num_nodes <= 0 || num_nodes > 0x7fffffffL) { set_msg(msg, msg_len, "num_nodes must be a positive 32-bit integer.\n"); return PROTON_EM_EINVAL; }

  // ---- analyical_wrapper: output directory, curden.csv, B.csv ------------------------------------
  const size_t pos = input_files.find("input");
  if (pos == std::string::npos) { set_msg(msg, msg_len, "The input directory path does not contain \"input\".\n"); return PROTON_EM_EINVAL; }
  std::string out_dir = input_files;
  out_dir.replace(pos, 5, "output");

  std::vector<double> u;
  {
    const std::string path = input_files + "curden.csv";
    std::ifstream f(path);
    if (!f) { set_msg(msg, msg_len, "The file " + path + " was not found.\nTerminating.\n"); return PROTON_EM_EIO; }
    std::string line;
    size_t lineno = 0;
    while (std::getline(f, line)) {
      ++lineno;
      char* end = nullptr;
      errno = 0;
      const double v = std::strtod(line.c_str(), &end);
      if (end == line.c_str()) {
        set_msg(msg, msg_len, "Error parsing file " + path + " at line " + std::to_string(lineno) + ".\nTerminating.\n");
        return PROTON_EM_EINVAL;
      }
      u.push_back(v);
    }
  }
  std::vector<int32_t> rows, cols;
  std::vector<double> vals;
  {
    const std::string path = input_files + "B.csv";
    std::ifstream f(path);
    if (!f) { set_msg(msg, msg_len, "Terminating\nThe file " + path + " was not found.\nTerminating.\n"); return PROTON_EM_EIO; }
    std::string line;
    size_t lineno = 0;
    while (std::getline(f, line)) {
      ++lineno;
      const std::vector<std::string> tok = tokenize(line);
      if (tok.empty()) continue;
      if (tok.size() < 2) { set_msg(msg, msg_len, "Required second value at line " + std::to_string(lineno) + " of the file " + path + ".\nTerminating.\n"); return PROTON_EM_EINVAL; }
      if (tok.size() < 3) { set_msg(msg, msg_len, "Required third value at line " + std::to_string(lineno) + " of the file " + path + ".\nTerminating.\n"); return PROTON_EM_EINVAL; }
      char* e1 = nullptr; char* e2 = nullptr;
      const long r = std::strtol(tok[0].c_str(), &e1, 10), c = std::strtol(tok[1].c_str(), &e2, 10);
      if (e1 == tok[0].c_str() || e2 == tok[1].c_str()) {
        set_msg(msg, msg_len, "Error parsing file " + path + " at line " + std::to_string(lineno) + ".\nTerminating.\n");
        return PROTON_EM_EINVAL;
      }
      rows.push_back((int32_t)r); cols.push_back((int32_t)c); vals.push_back(std::atof(tok[2].c_str()));
    }
  }
  for (size_t i = 0; i < rows.size(); ++i) {
    if (rows[i] < 0 || rows[i] >= num_nodes || cols[i] < 0 || (size_t)cols[i] >= u.size()) {
      set_msg(msg, msg_len, "B.csv entry " + std::to_string(i + 1) + " lies outside the " + std::to_string(num_nodes) + " x " +
                                std::to_string(u.size()) + " matrix.\nTerminating.\n");
      return PROTON_EM_EINVAL;
    }
  }

  // ---- analytical(): one line, one timestamp on the GPU -----------------------------------------
  if (!ctx) { set_msg(msg, msg_len, "No CUDA context.\n"); return PROTON_EM_ECUDA; }
  const int32_t N = (int32_t)num_nodes;
  const int64_t trip_off[2] = {0, (int64_t)rows.size()};
  const int64_t u_off[2] = {0, (int64_t)u.size()};
  proton_em_batch batch{1, &N, &a_coeff, trip_off, rows.data(), cols.data(), vals.data(), u_off, u.data()};
  std::vector<double> stress((size_t)N);
  double mx = 0.0; int32_t amx = 0; uint8_t flag = 0;
  proton_em_result res{stress.data(), &mx, &amx, &flag};
  proton_em_opts opts{0.0, PROTON_EM_ALG_AUTO, 0};
  const int rc = proton_em_analyze_host(ctx, &batch, &sim_time, 1, &opts, &res, 0);
  if (rc != PROTON_EM_OK) { set_msg(msg, msg_len, std::string(proton_em_last_error()) + "\nTerminating.\n"); return rc; }

  // ---- stress_<%f>.txt, one "%.18g" per line ----------------------------------------------------
  const std::string fname = out_dir + "stress_" + std::to_string(sim_time) + ".txt";
  const size_t slash = fname.find_last_of('/');
  if (slash != std::string::npos && !make_dirs(fname.substr(0, slash))) {
    set_msg(msg, msg_len, "Cannot create the output directory " + fname.substr(0, slash) + ".\nTerminating.\n");
    return PROTON_EM_EIO;
  }
  FILE* fo = std::fopen(fname.c_str(), "w");
  if (!fo) { set_msg(msg, msg_len, "Cannot open " + fname + " for writing.\nTerminating.\n"); return PROTON_EM_EIO; }
  for (int32_t i = 0; i < N; ++i) std::fprintf(fo, "%.18g\n", stress[(size_t)i]);
  std::fclose(fo);
  return PROTON_EM_OK;
}
