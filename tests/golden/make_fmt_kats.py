#!/usr/bin/env python
"""Known-answer vectors of the reference's number formatting: compiles a tiny program against the fmt the reference vendors
(/root/reference/fmt/include, header-only mode) and records fmt::format("{:.Nf}", x) for a sweep of values -- ordinary ones,
and the tiny non-zero ones for which fmt 5.3.1 prints an extra digit.  Run HERE; output: tests/golden/fmt_fixed_kats.json."""
import json
import os
import random
import subprocess
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = r'''
#define FMT_HEADER_ONLY
#include <fmt/format.h>
#include <cstdio>
#include <cstdlib>
int main() { char hex[64]; int p; while (scanf("%63s %d", hex, &p) == 2) { double x = strtod(hex, 0);
  std::string s = p == 0 ? fmt::format("{:.0f}", x) : p == 2 ? fmt::format("{:.2f}", x) : p == 3 ? fmt::format("{:.3f}", x) : fmt::format("{:.6f}", x);
  printf("%s\n", s.c_str()); } }
'''


def main():
    rng = random.Random(20240607)
    vals = [0.0, -0.0, 1.0, 0.5, 0.0005, 0.00049999, 1e-4, 9.9999e-5, 9.6e-5, 5e-5, 4.9e-5, 1e-7, 10000.0, 5000.0, 12.3456789,
            0.9995, 0.9994999, 2.5e-3, 4.3429448190325176e-16, 9.643274665402793e-16, 1e-300, 2.2250738585072014e-308, 5e-324,
            float("nan"), float("inf"), -float("inf")]
    for e in range(-20, 3):
        for _ in range(40):
            vals.append(rng.uniform(1.0, 10.0) * 10.0 ** e)
        for m in (1.0, 2.0, 4.9999, 5.0, 5.0001, 9.99):
            vals.append(m * 10.0 ** e)
    vals += [-v for v in vals[20:200:7]]
    cases = [(v, p) for v in vals for p in (0, 2, 3, 6)]
    d = tempfile.mkdtemp()
    with open(os.path.join(d, "k.cpp"), "w") as f:
        f.write(SRC)
    subprocess.run(["g++", "-std=c++11", "-O1", "-I/root/reference/fmt/include", "k.cpp", "-o", "k"], cwd=d, check=True)
    inp = "".join("%s %d\n" % (float(v).hex() if v == v and abs(v) != float("inf") else repr(float(v)), p) for v, p in cases)
    out = subprocess.run([os.path.join(d, "k")], input=inp, capture_output=True, text=True, check=True).stdout.split("\n")
    kats = [[float(v).hex() if v == v and abs(v) != float("inf") else repr(float(v)), p, out[i]] for i, (v, p) in enumerate(cases)]
    with open(os.path.join(HERE, "fmt_fixed_kats.json"), "w") as f:
        json.dump(kats, f, separators=(",", ":"))
    quirk = sum(1 for h, p, s in kats if "n" not in h and s != ("%.*f" % (p, float.fromhex(h))))
    print(len(kats), "vectors,", quirk, "differ from printf")


if __name__ == "__main__":
    main()
