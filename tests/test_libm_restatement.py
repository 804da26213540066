"""The path tracer must reproduce the reference's CPU build bit for bit, and that build calls glibc's sinf / cosf / acosf, none of
which is correctly rounded. pt_device.cuh therefore restates glibc's own algorithms for the device (glibcSinCosf, glibcAcosf).
This test compiles exactly those blocks of the header for the host (CUDA intrinsics shimmed) and compares them with the libm of
the machine over the whole argument range the kernels use: every 61st float of [0, 2 pi] resp. [-1, 1] plus dense neighbourhoods
of the range breaks (the exhaustive sweep, 1.09e9 + 2.13e9 arguments, was run once while writing them: 0 mismatches)."""
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

HARNESS = r"""
#include <math.h>
#include <stdio.h>
#include <stdint.h>
#include <string.h>
#define __device__
#define __forceinline__ inline
static inline uint32_t __float_as_uint(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static inline int32_t __float_as_int(float f) { int32_t u; memcpy(&u, &f, 4); return u; }
static inline float __int_as_float(int32_t i) { float f; memcpy(&f, &i, 4); return f; }
static inline int __double2int_rz(double d) { return (int)d; }
%s
static inline float fromBits(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
static long bad = 0, total = 0;
static void checkSinCos(uint32_t u) {
  const float x = fromBits(u); float s, c; glibcSinCosf(x, s, c); const float ws = sinf(x), wc = cosf(x); total++;
  if (memcmp(&s, &ws, 4) || memcmp(&c, &wc, 4)) { if (bad < 5) printf("sincos %%a: %%a %%a vs %%a %%a\n", x, s, c, ws, wc); bad++; }
}
static void checkAcos(uint32_t u) {
  const float x = fromBits(u); const float a = glibcAcosf(x), w = acosf(x); total++;
  if (memcmp(&a, &w, 4) && !(a != a && w != w)) { if (bad < 5) printf("acos %%a: %%a vs %%a\n", x, a, w); bad++; }
}
int main() {
  const uint32_t twoPi = 0x40C90FDBu;                     // 2 pi rounded to float: the largest argument of the diffuse sampler
  for (uint32_t u = 0; u <= twoPi + 8; u += 61) checkSinCos(u);
  const uint32_t breaks[] = {0x39800000u /* 2^-12 */, 0x3f490fdbu /* pi/4 */, 0x3fc90fdbu /* pi/2 */, 0x4016cbe4u /* 3pi/4 */, 0x40490fdbu /* pi */,
                             0x407b53d1u /* 5pi/4 */, 0x4096cbe4u /* 3pi/2 */, 0x40afede0u /* 7pi/4 */, twoPi};
  for (uint32_t b : breaks) for (uint32_t u = b - 20000; u <= b + 20000; ++u) checkSinCos(u);
  for (uint32_t u = 0; u <= 0x3f800000u + 4; u += 61) { checkAcos(u); checkAcos(u | 0x80000000u); }
  const uint32_t abreaks[] = {0x32800000u /* 2^-26 */, 0x3f000000u /* 0.5 */, 0x3f800000u /* 1 */};
  for (uint32_t b : abreaks) for (uint32_t u = b - 20000; u <= b + 20000; ++u) { checkAcos(u); checkAcos(u | 0x80000000u); }
  printf("checked %%ld mismatches %%ld\n", total, bad);
  return bad ? 1 : 0;
}
"""


def test_device_sinf_cosf_acosf_are_the_hosts_libm(tmp_path):
    src = open(os.path.join(ROOT, "ipu_path_trace_b200", "csrc", "pt_device.cuh")).read()
    blocks = re.findall(r"// \[host-checkable:[^\n]*\n(.*?)// \[end host-checkable\]", src, flags=re.S)
    assert len(blocks) == 2 and "glibcSinCosf" in blocks[0] and "glibcAcosf" in blocks[1]
    cpp = tmp_path / "libm_check.cpp"
    cpp.write_text(HARNESS % "\n".join(blocks))
    exe = tmp_path / "libm_check"
    # no FMA contraction and no builtin folding: the device build uses -fmad=false, and the comparison must call the real libm
    subprocess.run(["/usr/bin/g++", "-O2", "-fno-builtin", "-ffp-contract=off", str(cpp), "-o", str(exe), "-lm"], check=True)
    r = subprocess.run([str(exe)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:]
    assert "mismatches 0" in r.stdout and int(r.stdout.split()[1]) > 5 * 10 ** 7
