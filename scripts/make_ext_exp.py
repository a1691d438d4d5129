"""timing-experiment copies of csrc/cdx_query.cu (-> csrc/build/exp, git-ignored; results are wrong by design, the product source
carries no experiment switches).  Ablations of the warp-specialised extended query gpis_extw_kernel:
  1  generators store the coordinate difference instead of evaluating k*      (cost of the generation arithmetic)
  2  epilogue uses the coordinate differences instead of evaluating k, c      (cost of the epilogue arithmetic)
  3  consumers issue no fragment loads and no DMMAs                           (everything but the tensor work)
Usage: python scripts/make_ext_exp.py 1 2 3   then   python scripts/ext_exp.py <n>   (DESIGN.md section 4 quotes the numbers;
profiles/r02_ext_ablations.json holds the runs of both the single-role kernel and this one)."""
import os
import subprocess
import sys

root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
csrc = os.path.join(root, "compliancedex_b200", "csrc")
src = open(os.path.join(csrc, "cdx_query.cu")).read()
a = src.index("gpis_extw_kernel(const double* __restrict__ Xq")
head, body = src[:a], src[a:]

GEN = "        cdx_kernel_k_batch<KERNEL, QPG>(sq, kp, kv);\n"
EPI = "            cdx_kernel_eval<KERNEL>(fma(dx, dx, fma(dy, dy, dz * dz)), kp, k, cc);\n            const double w = acc[i][j][e];"
MMA = "#pragma unroll\n      for (int s4 = 0; s4 < 4; s4++) {\n        const int xo = ((2 * s4 + xs) ^ xr) * 2;"
for pat in (GEN, EPI, MMA):
    assert body.count(pat) == 1, pat
body = body.replace(GEN, "#if EXP == 1\n        for (int i = 0; i < QPG; i++) kv[i] = sq[i];\n#else\n" + GEN + "#endif\n")
body = body.replace(EPI, "#if EXP == 2\n            k = dx; cc = dy + dz;\n#else\n" + EPI.split("\n")[0] + "\n#endif\n" + EPI.split("\n")[1])
body = body.replace(MMA, "#if EXP == 3\n      if (c < 0)\n#endif\n" + MMA)
out = os.path.join(csrc, "build", "exp")
os.makedirs(out, exist_ok=True)
text = (head + body).replace('#include "cdx_internal.cuh"', '#include "../../cdx_internal.cuh"\n#ifndef EXP\n#define EXP 0\n#endif', 1)
open(os.path.join(out, "cdx_query_exp.cu"), "w").write(text)
objs = [os.path.join(csrc, "build", f"{n}.o") for n in ("cdx_fit", "cdx_gemm", "cdx_query_i8x", "cdx_grid", "cdx_pregrasp", "cdx_sdf")]
for e in sys.argv[1:]:
    o = os.path.join(out, f"cdx_query_exp{e}.o")
    subprocess.check_call(["nvcc", "-O3", "-lineinfo", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-Xcompiler", "-fPIC",
                           f"-DEXP={e}", "-I", os.path.join(root, "include"), "-c", os.path.join(out, "cdx_query_exp.cu"), "-o", o])
    subprocess.check_call(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", os.path.join(out, f"libcdx_ext_exp{e}.so"), o] + objs + ["-ldl"])
    print("built", e)
