"""timing-experiment copy of csrc/cdx_query_i8x.cu (-> csrc/build/exp, git-ignored): -DEXP=1 no k* arithmetic, 2 no MMA issued,
4 no arithmetic and no stores (MMA + operand streaming only), 7 generators idle half the FP64 (no Newton step).  Results are wrong
by design; the product source carries no experiment switches."""
import os
import re
import sys

root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
src = open(os.path.join(root, "compliancedex_b200/csrc/cdx_query_i8x.cu")).read()
s = src.replace('#include "cdx_internal.cuh"', '#include "../../cdx_internal.cuh"\n#ifndef EXP\n#define EXP 0\n#endif', 1)


def rep(a, b):
    global s
    assert a in s, a[:60]
    s = s.replace(a, b)


rep('''              const double2 a = P2[2 * j], b = P2[2 * j + 1];
              const double sd =''', '''              const double2 a = P2[2 * j], b = P2[2 * j + 1];
              if (EXP == 1 || EXP == 4) { L[p] = (unsigned)__double2loint(a.x); H[p] = (unsigned)__double2hiint(b.y); continue; }
              if (EXP == 7) {   // the same instruction sequence in FP32 on reinterpreted bits: no FP64-pipe instruction at all
                const float ax = __int_as_float(__double2hiint(a.x)), ay = __int_as_float(__double2hiint(a.y)),
                            bx = __int_as_float(__double2hiint(b.x)), by_ = __int_as_float(__double2hiint(b.y));
                const float xf = __int_as_float(__double2hiint(xq)), yf = __int_as_float(__double2hiint(yq)), zf = __int_as_float(__double2hiint(zq));
                const float sf = fmaf(xf, ax, fmaf(yf, ay, fmaf(zf, bx, by_ + 1.0f)));
                const float yy = xrsqrt_ftz(sf);
                float df = sf * yy;
                df = fmaf(fmaf(-df, df, sf), 0.5f * yy, df);
                const float rf = fmaf(sf, fmaf(2.0f, df, -3.0f), 12582912.0f);
                L[p] = __float_as_uint(rf); H[p] = __float_as_uint(df);
                continue;
              }
              const double sd =''')
rep('''                if (PAIR) xtc_mma_i8_pair(tmem + d * NQ, da, db, idesc, (c > 0 || i > 0) ? 1u : 0u);
                else xtc_mma_i8(tmem + d * NQ, da, db, idesc, (c > 0 || i > 0) ? 1u : 0u);''', '''                if (EXP == 2) continue;
                if (PAIR) xtc_mma_i8_pair(tmem + d * NQ, da, db, idesc, (c > 0 || i > 0) ? 1u : 0u);
                else xtc_mma_i8(tmem + d * NQ, da, db, idesc, (c > 0 || i > 0) ? 1u : 0u);''')
rep('''          if (active) {
#pragma unroll
            for (int i = 0; i < S; i++) xst_local16(''', '''          if (active && EXP != 4) {
#pragma unroll
            for (int i = 0; i < S; i++) xst_local16(''')
rep('''            if (C > 1 && !PAIR) xmb_arrive_expect_tx(b_full + 8 * s, (unsigned)(nact * S * 16 * (C - 1)));''',
    '''            if (C > 1 && !PAIR) xmb_arrive_expect_tx(b_full + 8 * s, EXP == 4 ? 0u : (unsigned)(nact * S * 16 * (C - 1)));''')
os.makedirs(os.path.join(root, "compliancedex_b200/csrc/build/exp"), exist_ok=True)
open(os.path.join(root, "compliancedex_b200/csrc/build/exp/cdx_query_i8x.cu"), "w").write(s)
print("written")
