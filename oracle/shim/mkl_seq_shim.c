/* oracle/shim/mkl_seq_shim.c — TEST INFRASTRUCTURE ONLY. Empty library: only its soname
 * (libmkl_sequential.so.2) is needed by the reference binary's DT_NEEDED list. */
int em_shim_mkl_sequential_present = 1;
