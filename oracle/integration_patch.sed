# oracle/integration_patch.sed -- the patch of INTEGRATION.md section 1 as an edit script: applied
# to a scratch copy of the reference's OWN main program (/root/reference/matFact.c, read where it
# lies) it replaces the two calls at the seam (matFact.c:124 and :127) by one call into the C ABI of
# include/matfact_b200.h.  `make -C oracle ref` builds oracle/_ref/matFact_patched from the result
# (the patched source is a temporary file and is deleted; nothing of the reference is stored in
# this repository).  tests/test_gpu_parity.py::test_reference_main_with_the_patch runs it.
1i #include "matfact_b200.h"
s|^\tmatrix_factorization(B, L, R, entries, non_zero, iters, alpha);$|\tint *mfgpu_best = malloc(sizeof(int) * (users > 0 ? users : 1));\n\tif (mfgpu_run(users, items, features, non_zero, (const mfgpu_entry *)entries, iters, alpha, mat2d_data(L), mat2d_data(R), mfgpu_best, 1))\n\t\tdie((char *)mfgpu_last_error());|
s|^\tprint_output(B, entries);$|\tfor (int mfgpu_i = 0; mfgpu_i < users; mfgpu_i++)\n\t\tif (mfgpu_best[mfgpu_i] != -1)\n\t\t\tprintf("%d\\n", mfgpu_best[mfgpu_i]);\n\tfree(mfgpu_best);|
