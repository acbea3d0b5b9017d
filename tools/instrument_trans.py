"""Development aid: build a copy of the library whose transition kernel (k_solve<.., ROUND = true>) prints, for chosen rounds,
the clock64 cycles its lone/last tiles spend in the prologue (state restore), in every state handler, inside the L-BFGS
two-loop recursion, and in the epilogue.  This is how the figures quoted in csrc/mpcb_solve.cuh (lbfgs_apply) and
profiles/r02_summary.md section 6 were measured.  Usage: python tools/instrument_trans.py  ->  build/libmpcb200_tt.so, then
MPCB_LIB=$PWD/build/libmpcb200_tt.so MPCB_TT_ROUNDS=60:72 PROBE=rounds:8:1 python tools/mode_probe.py | grep '^T rnd'."""
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "model-predictive-control_b200", "csrc")
DST = os.path.join(ROOT, "build", "csrc_tt")


def rep(s, old, new):
    assert s.count(old) == 1, (s.count(old), old[:70])
    return s.replace(old, new)


def main():
    shutil.rmtree(DST, ignore_errors=True)
    shutil.copytree(SRC, DST)
    p = os.path.join(DST, "mpcb_solve.cuh")
    s = open(p).read()
    s = rep(s, "  extern __shared__ double smem[];\n  const int n = d.n, m = d.m, mem = sp.lbfgs_mem;",
            "  extern __shared__ double smem[];\n  const long long tk0 = clock64();\n  long long tseg[16] = {0}; long long tlast = 0; int slast = 15;\n"
            "  const int n = d.n, m = d.m, mem = sp.lbfgs_mem;")
    s = rep(s, "  for (;;) {\n    // ================= transition phase: run the tile's state machine up to its next evaluation ====\n    want = 0;\n    for (; !dead;) {",
            "  const long long tk1 = clock64(); tlast = tk1;\n  for (;;) {\n    want = 0;\n    for (; !dead;) {\n"
            "      { const long long tc = clock64(); tseg[slast] += tc - tlast; tlast = tc; slast = state; }")
    s = rep(s, "    if constexpr (ROUND) {\n      // ================= round mode: publish the request, save the registers, leave ===============",
            "    const long long tk2 = clock64(); tseg[slast] += tk2 - tlast;\n    if constexpr (ROUND) {")
    s = rep(s, "          sdb[(r++) * MPCB_LB] = __longlong_as_double((long long)b);\n        }\n      }\n      return;",
            "          sdb[(r++) * MPCB_LB] = __longlong_as_double((long long)b);\n        }\n      }\n"
            "      if (!notile && l == 0 && rnd >= d.tt_lo && rnd < d.tt_hi && (a.ctl == nullptr || a.ctl[0] <= 2u || a.nt == 1))\n"
            "        printf(\"T rnd %d tile %lld state->%d pairs %d: restore %lld machine %lld exit %lld | top %lld fd %lld lss %lld lsstep %lld other %lld\\n\",\n"
            "               (int)rnd, (long long)tile, state, lb_full ? mem : lb_idx, tk1 - tk0, tk2 - tk1, clock64() - tk2, tseg[6], tseg[7], tseg[8], tseg[9],\n"
            "               tseg[0] + tseg[1] + tseg[2] + tseg[3] + tseg[4] + tseg[5] + tseg[10] + tseg[11] + tseg[12]);\n"
            "      return;")
    open(p, "w").write(s)
    # the round window comes in through two spare KDesc fields filled from the environment
    p = os.path.join(DST, "mpcb_device.cuh")
    s = open(p).read()
    s = rep(s, "  int exact;", "  int exact; int tt_lo, tt_hi;")
    open(p, "w").write(s)
    p = os.path.join(DST, "mpcb_api.cu")
    s = open(p).read()
    s = rep(s, "  h->kd.debug = h->debug_timers;",
            "  h->kd.debug = h->debug_timers;\n  { const char* e = std::getenv(\"MPCB_TT_ROUNDS\"); int lo = 60, hi = 72; if (e) std::sscanf(e, \"%d:%d\", &lo, &hi); h->kd.tt_lo = lo; h->kd.tt_hi = hi; }")
    open(p, "w").write(s)
    out = os.path.join(ROOT, "build", "libmpcb200_tt.so")
    cmd = ["/usr/local/cuda/bin/nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-fmad=false",
           "-Xcompiler", "-fPIC", "-shared", "-I", os.path.join(ROOT, "include"), "-o", out, os.path.join(DST, "mpcb_api.cu"), "-ccbin", "/usr/bin/g++"]
    subprocess.check_call(cmd)
    print(out)


if __name__ == "__main__":
    sys.exit(main())
