/* oracle/shim/mkl_core_shim.c — TEST INFRASTRUCTURE ONLY. Thread-count setters the reference
 * binary calls from main (bin/EMtool_analytical@0x6964-0x698e); no-ops here. */
void MKL_Set_Num_Threads(int n) { (void)n; }
int MKL_Set_Num_Threads_Local(int n) { (void)n; return 1; }
