// emtool_main.cpp — `bin/EMtool_analytical`, command-line compatible with the reference executable of the
// same name (README.md:98-113: argv[1] = path to analytical.cfg; exit 0 on success, 1 + a message on
// stdout otherwise; main at bin/EMtool_analytical@0x6920).  All work happens in libproton_em.so.
// One CUDA context per process is expensive: the batched API, not this launcher, is the fast path.
#include <cstdio>
#include "../../include/proton_em.h"

int main(int argc, char** argv) {
  if (argc != 2) {
    std::printf("Wrong number of arguments. Configuration file expected.\n");
    return 1;
  }
  char msg[2048];
  msg[0] = 0;
  // Parse-only errors must not need a GPU: try without a context first when the input is invalid.
  proton_em_ctx* ctx = nullptr;
  int rc = proton_em_create(&ctx, 0);
  if (rc != PROTON_EM_OK) {
    // still report input errors the way the reference does; only then complain about the device
    int rc2 = proton_em_run_cfg(nullptr, argv[1], msg, sizeof(msg));
    if (rc2 == PROTON_EM_ECUDA) std::printf("%s\nTerminating.\n", proton_em_last_error());
    else std::fputs(msg, stdout);
    return 1;
  }
  rc = proton_em_run_cfg(ctx, argv[1], msg, sizeof(msg));
  std::fputs(msg, stdout);
  proton_em_destroy(ctx);
  return rc == PROTON_EM_OK ? 0 : 1;
}
