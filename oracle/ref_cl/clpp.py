"""TEST INFRASTRUCTURE — makes the reference's OpenCL C sources parseable by g++.

Reads every *.cl, geom.h and bxdf_types.h under $FLUCTUS_REF/src and writes a rewritten copy into the
(git-ignored) output directory.  The rewrite is purely lexical and touches only what C++ cannot parse:

  (float3)(a, b, c)   ->  float3{a, b, c}           vector literals (a comma expression in C++); a braced
                                                    initialiser keeps the components' evaluation order
                                                    left to right, as clang-based OpenCL compilers do
                                                    (matters at ggx.cl:27: two rand(seed) in one literal)
  expr.xyz / expr.xy  ->  expr.xyz_() / expr.xy_()  swizzles (member calls on the shim's vector types)

  atomic_add_float3(ptr, v)  ->  atomic_add_float3((volatile float3 *)ptr, v)   (utils.cl:324,333: the
                      float* -> float3* pointer conversion is implicit in C, an error in C++)

Every other token — arithmetic, operand order, control flow, macros, #ifdef structure — is the
reference's.  Nothing produced here is committed; the recipe is.
"""
import os
import re
import sys

VEC = re.compile(r"\(\s*(v?)(float[234]|int2)\s*\)\s*\(")
SWZ = re.compile(r"\.(xyz|xy)\b(?!\s*\()")
PTR = re.compile(r"return atomic_add_float([34])\(ptr,")


def _vector_literals(text):
    """Rightmost literal first, so that nested literals are rewritten inside out."""
    for m in reversed(list(VEC.finditer(text))):
        depth, k = 1, m.end()
        while depth:
            c = text[k]
            depth += (c == "(") - (c == ")")
            k += 1
        text = text[:m.start()] + m.group(1) + m.group(2) + "{" + text[m.end():k - 1] + "}" + text[k:]
    return text


def rewrite(text):
    text = _vector_literals(text)
    text = PTR.sub(lambda m: "return atomic_add_float%s((volatile float%s *)ptr," % (m.group(1), m.group(1)), text)
    return SWZ.sub(lambda m: ".%s_()" % m.group(1), text)


def main(src, dst):
    os.makedirs(dst, exist_ok=True)
    n = 0
    for name in sorted(os.listdir(src)):
        if not (name.endswith(".cl") or name in ("geom.h", "bxdf_types.h")):
            continue
        with open(os.path.join(src, name), "r", errors="replace") as f:
            out = rewrite(f.read())
        path = os.path.join(dst, name)
        if not os.path.exists(path) or open(path).read() != out:
            with open(path, "w") as f:
                f.write(out)
        n += 1
    return n


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
