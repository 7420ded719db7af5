"""TEST INFRASTRUCTURE ONLY.  Rewrites CUDA kernel launches `kernel<T><<<grid, block, smem, stream>>>(args);` into
`cuda_emu::LaunchKernel(grid, block, [&](auto&&... a) { kernel<T>(a...); }, args);` so that a .cu / .cuh source of
or-tools_b200/csrc compiles with g++ against tests/emu_lib/include (see cuda_runtime.h there).  Nothing else in the
source is touched; the line count is preserved so that compiler messages point at the original lines.

    python rewrite_launches.py <source> <destination>
"""
import sys


def _matching(text, start, open_ch, close_ch):
    depth = 0
    for i in range(start, len(text)):
        if text[i] == open_ch:
            depth += 1
        elif text[i] == close_ch:
            depth -= 1
            if depth == 0:
                return i
    raise ValueError("unbalanced " + open_ch)


def _name_start(text, end):
    """Start of the kernel name (identifier, optionally followed by balanced <...>) that ends at `end` (exclusive)."""
    i = end
    if text[i - 1] == ">":
        depth = 0
        while True:
            i -= 1
            if text[i] == ">":
                depth += 1
            elif text[i] == "<":
                depth -= 1
                if depth == 0:
                    break
    while i > 0 and (text[i - 1].isalnum() or text[i - 1] in "_:"):
        i -= 1
    return i


def _split_top_level(text):
    parts, depth, cur = [], 0, ""
    for ch in text:
        if ch in "(<[{":
            depth += 1
        elif ch in ")>]}":
            depth -= 1
        if ch == "," and depth == 0:
            parts.append(cur)
            cur = ""
        else:
            cur += ch
    parts.append(cur)
    return [p.strip() for p in parts]


def rewrite(text):
    out, pos, count = "", 0, 0
    while True:
        k = text.find("<<<", pos)
        if k < 0:
            return out + text[pos:], count
        close = text.index(">>>", k)
        name_at = _name_start(text, k)
        name = text[name_at:k]
        config = _split_top_level(text[k + 3:close])
        if len(config) < 2:
            raise ValueError("launch configuration with fewer than two entries: " + text[k:close + 3])
        paren = text.index("(", close)
        if text[close + 3:paren].strip():
            raise ValueError("unexpected text between >>> and (: " + text[close:paren + 1])
        end = _matching(text, paren, "(", ")")
        args = text[paren + 1:end]
        call = f"cuda_emu::LaunchKernel({config[0]}, {config[1]}, [&](auto&&... emu_args) {{ {name}(emu_args...); }}"
        if args.strip():
            call += ", " + args
        out += text[pos:name_at] + call + ")"
        pos = end + 1
        count += 1


if __name__ == "__main__":
    src, dst = sys.argv[1], sys.argv[2]
    new, n = rewrite(open(src).read())
    open(dst, "w").write(new)
    print(f"{src}: {n} launches rewritten")
