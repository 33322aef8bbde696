"""One-off source transformation (kept for the record of what was done to the kernel files):
  * every ``__global__`` kernel DEFINITION gets ``vb::pdl_sync();`` as its first statement (griddepcontrol.wait,
    then griddepcontrol.launch_dependents): a no-op for ordinary launches, the dependency point of a launch made
    with programmatic stream serialisation;
  * every ``kernel<<<grid, block, smem, stream>>>(args)`` becomes ``vb::launch_k(kernel, grid, block, smem, stream,
    args)`` (common.cuh), which adds that launch attribute when vb_set_pdl(1) is in force.
Usage: python scripts/pdl_transform.py file.cu [file.cu ...]   (idempotent)"""
import re
import sys


def match_paren(s, i, open_c="(", close_c=")"):
    assert s[i] == open_c, (s[i - 20:i + 20])
    depth = 0
    while i < len(s):
        c = s[i]
        if c == open_c:
            depth += 1
        elif c == close_c:
            depth -= 1
            if depth == 0:
                return i
        i += 1
    raise ValueError("unbalanced")


def split_top(s):
    out, depth, cur = [], 0, ""
    for c in s:
        if c in "(<[{":
            depth += 1
        elif c in ")>]}":
            depth -= 1
        if c == "," and depth == 0:
            out.append(cur.strip())
            cur = ""
        else:
            cur += c
    out.append(cur.strip())
    return out


def add_prologues(src):
    out, pos, n = "", 0, 0
    for m in re.finditer(r"__global__", src):
        i = m.end()
        # skip to the parameter list: first identifier followed by "(" that is not __launch_bounds__
        while True:
            mm = re.compile(r"([A-Za-z_]\w*)\s*\(").search(src, i)
            if mm.group(1) == "__launch_bounds__":
                i = match_paren(src, mm.end() - 1) + 1
                continue
            break
        close = match_paren(src, mm.end() - 1)
        j = close + 1
        while src[j] in " \t\r\n\\":
            j += 1
        if src[j] != "{":
            continue                                   # a declaration
        body = src[j + 1:j + 80]
        if "pdl_sync" in body:
            continue
        out += src[pos:j + 1] + "\n  vb::pdl_sync();"
        pos = j + 1
        n += 1
    return out + src[pos:], n


def convert_launches(src):
    out, pos, n = "", 0, 0
    while True:
        k = src.find("<<<", pos)
        if k < 0:
            break
        # kernel expression: walk back over an optional <...> and the identifier
        e = k
        b = e
        if src[b - 1] == ">":
            depth, b = 0, b - 1
            while True:
                if src[b] == ">":
                    depth += 1
                elif src[b] == "<":
                    depth -= 1
                    if depth == 0:
                        break
                b -= 1
        while b > 0 and (src[b - 1].isalnum() or src[b - 1] in "_:"):
            b -= 1
        kern = src[b:e]
        c_end = src.find(">>>", k)
        cfg = split_top(src[k + 3:c_end])
        while len(cfg) < 4:
            cfg.append("0")
        a0 = c_end + 3
        while src[a0] in " \t\r\n\\":
            a0 += 1
        a1 = match_paren(src, a0)
        args = src[a0 + 1:a1].strip()
        call = "vb::launch_k(%s, dim3(%s), dim3(%s), %s, %s%s)" % (kern, cfg[0], cfg[1], cfg[2], cfg[3],
                                                                  (", " + args) if args else "")
        out += src[pos:b] + call
        pos = a1 + 1
        n += 1
    return out + src[pos:], n


if __name__ == "__main__":
    for path in sys.argv[1:]:
        s = open(path).read()
        s, n1 = add_prologues(s)
        s, n2 = convert_launches(s)
        open(path, "w").write(s)
        print("%s: %d kernel prologues, %d launches" % (path, n1, n2))
