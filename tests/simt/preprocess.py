"""Rewrite the two pieces of CUDA syntax a host C++ compiler cannot parse, nothing else:

    kernel<<<grid, block, smem, stream>>>(args);
        -> simt::launch(dim3(grid), dim3(block), smem, [&]() { kernel(args); });
    extern __shared__ __align__(16) T name[];
        -> T* name = reinterpret_cast<T*>(simt::dyn_smem());

    python tests/simt/preprocess.py csrc/query.cu build/simt/query.cpp
"""
from __future__ import annotations

import re
import sys


def split_top_level(text: str):
    parts, depth, cur = [], 0, []
    for ch in text:
        if ch in "([{":
            depth += 1
        elif ch in ")]}":
            depth -= 1
        if ch == "," and depth == 0:
            parts.append("".join(cur).strip())
            cur = []
        else:
            cur.append(ch)
    parts.append("".join(cur).strip())
    return parts


def rewrite_launches(src: str) -> str:
    out, pos = [], 0
    name_re = re.compile(r"([A-Za-z_][\w:]*(?:\s*<[^<>;(){}]*>)?)\s*$")
    while True:
        i = src.find("<<<", pos)
        if i < 0:
            out.append(src[pos:])
            break
        m = name_re.search(src[pos:i])
        if not m:
            raise SystemExit(f"cannot find the kernel name before <<< at offset {i}")
        name_start = pos + m.start(1)
        j = src.find(">>>", i)
        config = split_top_level(src[i + 3:j])
        while len(config) < 4:
            config.append("0")
        k = j + 3
        while src[k].isspace():
            k += 1
        if src[k] != "(":
            raise SystemExit(f"expected '(' after >>> at offset {k}")
        depth, e = 0, k
        while True:
            depth += src[e] == "("
            depth -= src[e] == ")"
            if depth == 0:
                break
            e += 1
        args = src[k + 1:e]
        out.append(src[pos:name_start])
        out.append(f"simt::launch(\"{m.group(1)}\", dim3((unsigned)({config[0]})), dim3((unsigned)({config[1]})), "
                   f"(size_t)({config[2]}), [&]() {{ {m.group(1)}({args}); }})")
        pos = e + 1
    return "".join(out)


def rewrite_dynamic_smem(src: str) -> str:
    pat = re.compile(r"extern\s+__shared__\s+(?:__align__\(\d+\)\s+)?([A-Za-z_][\w ]*?)\s+(\w+)\[\];")
    return pat.sub(lambda m: f"{m.group(1)}* {m.group(2)} = reinterpret_cast<{m.group(1)}*>(simt::dyn_smem());", src)


def main() -> int:
    src = open(sys.argv[1]).read()
    src = rewrite_dynamic_smem(rewrite_launches(src))
    open(sys.argv[2], "w").write(f'#line 1 "{sys.argv[1]}"\n' + src)
    return 0


if __name__ == "__main__":
    sys.exit(main())
