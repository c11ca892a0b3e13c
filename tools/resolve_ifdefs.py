"""Development aid: resolves `#if MACRO` / `#elif` / `#else` arms of the given macros to fixed values in a source file
(a minimal unifdef).  Only conditions that are a bare macro name, or `A || B` of resolved names, are touched; the
`#ifndef MACRO / #define MACRO v / #endif` default blocks of resolved macros are dropped.
usage: resolve_ifdefs.py file MACRO=0 MACRO2=1 ...   (rewrites the file in place)"""
import re
import sys


def evaluate(cond, values):
    """True / False when every identifier of the condition is resolved, else None."""
    names = re.findall(r"[A-Za-z_]\w*", cond)
    if not names or any(n not in values for n in names):
        return None
    expr = re.sub(r"[A-Za-z_]\w*", lambda m: str(values[m.group(0)]), cond)
    expr = expr.replace("||", " or ").replace("&&", " and ").replace("!", " not ")
    return bool(eval(expr))


def main(path, values):
    lines = open(path).read().split("\n")
    out = []
    # stack entries: [resolved?, emitting_now, any_arm_taken, parent_emitting]
    stack = []
    i = 0
    emitting = True
    while i < len(lines):
        line = lines[i]
        s = line.strip()
        m = re.match(r"#\s*ifndef\s+(\w+)", s)
        if m and m.group(1) in values and i + 2 < len(lines) and re.match(r"#\s*define\s+" + m.group(1) + r"\b", lines[i + 1].strip()):
            # default block of a resolved macro: drop through its #endif
            j = i + 1
            while not re.match(r"#\s*endif", lines[j].strip()):
                j += 1
            i = j + 1
            continue
        m = re.match(r"#\s*if\s+(.*)$", s)
        if m and not s.startswith("#ifdef") and not s.startswith("#ifndef"):
            v = evaluate(m.group(1).split("//")[0].strip(), values)
            if v is None:
                stack.append([False, emitting, False, emitting])
                if emitting:
                    out.append(line)
            else:
                stack.append([True, emitting and v, v, emitting])
                emitting = emitting and v
            i += 1
            continue
        if re.match(r"#\s*(ifdef|ifndef)\b", s):
            stack.append([False, emitting, False, emitting])
            if emitting:
                out.append(line)
            i += 1
            continue
        m = re.match(r"#\s*elif\s+(.*)$", s)
        if m and stack and stack[-1][0]:
            top = stack[-1]
            v = evaluate(m.group(1).split("//")[0].strip(), values)
            if v is None:
                raise SystemExit("unresolvable #elif inside a resolved #if at line %d" % (i + 1))
            take = (not top[2]) and v
            top[2] = top[2] or v
            emitting = top[3] and take
            i += 1
            continue
        if re.match(r"#\s*else\b", s) and stack and stack[-1][0]:
            top = stack[-1]
            emitting = top[3] and not top[2]
            top[2] = True
            i += 1
            continue
        if re.match(r"#\s*endif\b", s):
            top = stack.pop()
            if top[0]:
                emitting = top[3]
            else:
                if emitting:
                    out.append(line)
            i += 1
            continue
        if emitting:
            out.append(line)
        i += 1
    open(path, "w").write("\n".join(out))


if __name__ == "__main__":
    main(sys.argv[1], {k: int(v) for k, v in (a.split("=") for a in sys.argv[2:])})
