#!/usr/bin/env python3
"""Machine-transcribe the reference's Generation I known-answer tests into Python fixtures.

Input : /root/reference/cpp/extern/engine/src/lib/gen1/test.zig (126 `test "..."` blocks; each scripts
        a battle with a fixed roll list, choices, expected HP deltas / state / results).
Output: tests/golden/gen1_kat_cases.py — one `def case_NNN(h)` per test, written against the API of
        tests/kat_harness.py (which mirrors the Zig harness, test.zig:10415-10554).

Only the Pokémon Showdown compatibility branch of every `if (showdown) A else B` is kept; protocol-log
expectations (`t.log.expected.*`) become calls on the harness's expected-log writer; probability expectations are dropped. Tests using constructs the
translator does not understand are emitted as skipped with the reason, never silently altered.

Run in the build container only (the reference tree is not on the GPU box); the output is committed.
"""
import os
import re
import sys

REF = "/root/reference/cpp/extern/engine/src/lib/gen1/test.zig"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class Untranslatable(Exception):
    pass


def strip_comments(src):
    out = []
    for line in src.splitlines():
        # no string literals containing // inside test bodies except test names (handled before)
        i = line.find("//")
        if i >= 0:
            line = line[:i]
        out.append(line)
    return "\n".join(out)


def match_close(s, i, open_ch, close_ch):
    """s[i] == open_ch; return index of the matching close."""
    depth = 0
    j = i
    while j < len(s):
        c = s[j]
        if c == '"':
            j = s.index('"', j + 1)
        elif c == open_ch:
            depth += 1
        elif c == close_ch:
            depth -= 1
            if depth == 0:
                return j
        j += 1
    raise Untranslatable("unbalanced " + open_ch)


def skip_ws(s, i):
    while i < len(s) and s[i].isspace():
        i += 1
    return i


def scan_statement(s, i):
    """Return index just past the `;` ending the simple statement starting at i."""
    depth = 0
    j = i
    while j < len(s):
        c = s[j]
        if c == '"':
            j = s.index('"', j + 1)
        elif c in "([{":
            depth += 1
        elif c in ")]}":
            depth -= 1
        elif c == ";" and depth == 0:
            return j + 1
        j += 1
    raise Untranslatable("unterminated statement")


KW = re.compile(r"(inline\s+for|for|while|if|switch)\b")


def parse_block(s, i, end):
    """Parse statements in s[i:end] into a list of nodes."""
    nodes = []
    while True:
        i = skip_ws(s, i)
        if i >= end:
            return nodes
        if s[i] == "{":
            j = match_close(s, i, "{", "}")
            nodes.append(("block", parse_block(s, i + 1, j)))
            i = j + 1
            continue
        m = KW.match(s, i)
        if m and m.group(1) == "if":
            i, node = parse_if(s, i, end)
            nodes.append(node)
            continue
        if m and m.group(1) in ("for", "inline for", "while", "switch") or (m and m.group(1).startswith("inline")):
            # header up to the block brace
            k = s.index("(", i)
            k = match_close(s, k, "(", ")") + 1
            k = skip_ws(s, k)
            if s[k] == "|":
                k = s.index("|", k + 1) + 1
            k = skip_ws(s, k)
            if s[k] == ":":  # while (...) : (...)
                k = s.index("(", k)
                k = match_close(s, k, "(", ")") + 1
                k = skip_ws(s, k)
            if s[k] != "{":
                se = scan_statement(s, k)
                nodes.append(("loop", s[i:k].strip(), [("stmt", s[k:se - 1].strip())]))
                i = se
                continue
            j = match_close(s, k, "{", "}")
            nodes.append(("loop", s[i:k].strip(), parse_block(s, k + 1, j)))
            i = j + 1
            continue
        j = scan_statement(s, i)
        nodes.append(("stmt", s[i:j - 1].strip()))
        i = j


def parse_if(s, i, end):
    k = s.index("(", i)
    ce = match_close(s, k, "(", ")")
    cond = s[k + 1:ce].strip()
    j = skip_ws(s, ce + 1)
    if s[j] == "|":
        raise Untranslatable("if with capture")
    if s[j] == "{":
        be = match_close(s, j, "{", "}")
        then = parse_block(s, j + 1, be)
        j = be + 1
    else:
        se = scan_statement(s, j)
        then = [("stmt", s[j:se - 1].strip())]
        j = se
    els = None
    k = skip_ws(s, j)
    if s.startswith("else", k) and not (s[k + 4].isalnum() or s[k + 4] == "_"):
        k = skip_ws(s, k + 4)
        if s.startswith("if", k) and not s[k + 2].isalnum():
            j, node = parse_if(s, k, end)
            els = [node]
        elif s[k] == "{":
            be = match_close(s, k, "{", "}")
            els = parse_block(s, k + 1, be)
            j = be + 1
        else:
            se = scan_statement(s, k)
            els = [("stmt", s[k:se - 1].strip())]
            j = se
    return j, ("if", cond, then, els)


# ----------------------------------------------------------------------------- expressions
def find_else(s, i):
    """From i (just after an if-condition) find the depth-0 `else` that pairs with it."""
    depth = 0
    pending = 0
    j = i
    while j < len(s):
        c = s[j]
        if c == '"':
            j = s.index('"', j + 1)
        elif c in "([{":
            depth += 1
        elif c in ")]}":
            if depth == 0:
                return -1
            depth -= 1
        elif depth == 0 and (c == ";" ):
            return -1
        elif depth == 0 and s.startswith("if", j) and (j == 0 or not (s[j - 1].isalnum() or s[j - 1] in "_.")) \
                and not (s[j + 2].isalnum() or s[j + 2] == "_"):
            pending += 1
        elif depth == 0 and s.startswith("else", j) and not (s[j - 1].isalnum() or s[j - 1] == "_") \
                and (j + 4 >= len(s) or not (s[j + 4].isalnum() or s[j + 4] == "_")):
            if pending == 0:
                return j
            pending -= 1
        j += 1
    return -1


def expr_end(s, i):
    """End of the expression starting at i: the first depth-0 `,` `)` `}` `]` `;` or end of string."""
    depth = 0
    j = i
    while j < len(s):
        c = s[j]
        if c == '"':
            j = s.index('"', j + 1)
        elif c in "([{":
            depth += 1
        elif c in ")]}":
            if depth == 0:
                return j
            depth -= 1
        elif c in ",;" and depth == 0:
            return j
        j += 1
    return len(s)


IF_SD = re.compile(r"(?<![\w.])if\s*\(\s*(!?)\s*showdown\s*\)")


def resolve_showdown(s):
    """Collapse every `if (showdown) A else B` expression to its showdown branch."""
    while True:
        m = IF_SD.search(s)
        if not m:
            return s
        neg = m.group(1) == "!"
        a0 = m.end()
        e = find_else(s, a0)
        if e < 0:
            raise Untranslatable("showdown-if expression without else: " + s[m.start():m.start() + 60])
        b0 = e + 4
        b1 = expr_end(s, b0)
        a, b = s[a0:e].strip(), s[b0:b1].strip()
        keep = b if neg else a
        s = s[:m.start()] + "(" + keep + ")" + s[b1:]


def convert_braces(s):
    """`.{...}`, `&.{...}`, `&[_]T{...}`, `Name{...}` -> python list / dict(...) / Name(...)."""
    out = []
    i = 0
    while i < len(s):
        c = s[i]
        if c == "{":
            j = match_close(s, i, "{", "}")
            inner = convert_braces(s[i + 1:j])
            # what precedes the brace?
            pre = "".join(out).rstrip()
            m_lit = re.search(r"(&?\.)$", pre)
            m_arr = re.search(r"(&?\[_\]\w+)$", pre)
            m_name = re.search(r"([A-Za-z_]\w*)$", pre)
            is_struct = re.match(r"\s*\.\w+\s*=", inner) is not None
            if is_struct:
                inner = re.sub(r"(^|,)\s*\.(\w+)\s*=", lambda mm: mm.group(1) + " " + pykey(mm.group(2)) + "=", inner)
            if m_lit:
                pre = pre[:m_lit.start()]
                if is_struct:
                    rep = "dict(" + inner + ")"
                elif inner.strip() == "":
                    rep = "EMPTY"
                else:
                    rep = "[" + inner + "]"
            elif m_arr:
                pre = pre[:m_arr.start()]
                rep = "[" + inner + "]"
            elif m_name and m_name.group(1) not in ("else", "try", "comptime", "return"):
                rep = "(" + inner + ")"
            else:
                raise Untranslatable("block brace inside expression: " + s[max(0, i - 30):i + 30])
            out = [pre, rep]
            i = j + 1
            continue
        out.append(c)
        i += 1
    return "".join(out)


def pykey(k):
    return {"def": "def_", "type": "type_"}.get(k, k)


def strip_builtin(s, name, repl):
    """@name(T, v) or @name(v) -> repl(args)."""
    while True:
        k = s.find("@" + name + "(")
        if k < 0:
            return s
        o = k + len(name) + 1
        c = match_close(s, o, "(", ")")
        args = split_args(s[o + 1:c])
        s = s[:k] + repl(args) + s[c + 1:]


def split_args(s):
    args, depth, cur = [], 0, []
    for ch in s:
        if ch in "([{":
            depth += 1
        elif ch in ")]}":
            depth -= 1
        if ch == "," and depth == 0:
            args.append("".join(cur).strip())
            cur = []
        else:
            cur.append(ch)
    if "".join(cur).strip():
        args.append("".join(cur).strip())
    return args


def translate_expr(s):
    s = resolve_showdown(s)
    s = re.sub(r"\bcomptime\s+", "", s)
    s = re.sub(r"\btry\s+", "", s)
    s = strip_builtin(s, "as", lambda a: "(" + a[1] + ")")
    s = strip_builtin(s, "intCast", lambda a: "(" + a[-1] + ")")
    s = strip_builtin(s, "intFromEnum", lambda a: "int(" + a[0] + ")")
    s = strip_builtin(s, "intFromBool", lambda a: "int(" + a[0] + ")")
    s = strip_builtin(s, "enumFromInt", lambda a: "(" + a[0] + ")")
    s = strip_builtin(s, "max", lambda a: "max(" + ", ".join(a) + ")")
    s = strip_builtin(s, "min", lambda a: "min(" + ", ".join(a) + ")")
    s = strip_builtin(s, "divTrunc", lambda a: "((" + a[0] + ") // (" + a[1] + "))")
    if "@" in s:
        raise Untranslatable("builtin: " + s[s.index("@"):s.index("@") + 30])
    s = convert_braces(s)
    s = re.sub(r"std\.math\.maxInt\(u8\)", "255", s)
    s = re.sub(r"std\.math\.maxInt\(u16\)", "65535", s)
    # slices
    s = re.sub(r"\[(\w+)\.\.(\w+)\]", r"[\1:\2]", s)
    # address-of
    s = re.sub(r"&(?=[A-Za-z_\[])", "", s)
    # field names that are python keywords
    s = re.sub(r"\.def\b", ".def_", s)
    s = re.sub(r"\.type\b(?!\d)", ".type_", s)
    # enum literals: a dot not preceded by an identifier char, `)` or `]`
    s = re.sub(r"(?<![\w)\]])\.([A-Za-z_]\w*)", r"L.\1", s)
    # operators
    s = re.sub(r"~\s*(\w+)", r"(MAX ^ \1)", s)
    s = re.sub(r"!(?!=)", " not ", s)
    s = s.replace("+%", "+").replace("-%", "-").replace("*%", "*")
    s = re.sub(r"(?<![/])/(?![/=])", "//", s)
    s = re.sub(r"\borelse\b", "or", s)
    s = re.sub(r"\btrue\b", "True", s)
    s = re.sub(r"\bfalse\b", "False", s)
    s = re.sub(r"\bnull\b", "None", s)
    s = re.sub(r"\bStatus\.is\(", "Status.is_(", s)
    s = s.replace("++", "+")
    s = re.sub(r"\.None\b", ".None_", s)
    s = re.sub(r"\bfrom\b", "from_", s)
    s = re.sub(r"\butil\.repeat\(", "repeat(", s)
    for bad in ("pkmn.", "rng.", "calc.", "helpers.", "std.", "chance."):
        if re.search(r"(?<![\w.])" + re.escape(bad), s):
            raise Untranslatable("uses " + bad + ": " + s[:60])
    return s.strip()


# ----------------------------------------------------------------------------- statements
SKIP_STMT = re.compile(
    r"^(try\s+)?(expected\.\w+\(|defer\b|try\s+expectLog\b|expectLog\b|"
    r"(var|const)\s+(expected_buf|actual_buf|expected_writer|actual_writer)\b|"
    r"(var|const)\s+(expected|actual)(:\s*FixedLog)?\s*=\s*(FixedLog|\.\{\s*\.writer)|"
    r"(var|const)\s+options\s*=\s*pkmn\.battle\.options|(var|const)\s+size\s*=|"
    r"expected_writer\.|actual_writer\.|try\s+t\.expectProbability|t\.expectProbability)")
DECL = re.compile(r"^(?:const|var)\s+(\w+)\s*(?::\s*[^=]+?)?\s*=\s*(.*)$", re.S)


def emit(nodes, ind, out):
    pad = "    " * ind
    n0 = len(out)
    for node in nodes:
        kind = node[0]
        if kind == "stmt":
            st = node[1]
            if not st or SKIP_STMT.match(st):
                continue
            if re.match(r"^return\s+error\.SkipZigTest$", st):
                out.append(pad + "raise SkipCase()")
                continue
            if re.match(r"^return\b", st):
                raise Untranslatable("return: " + st[:40])
            m = DECL.match(st)
            if m:
                out.append(pad + ("from_" if m.group(1) == "from" else m.group(1)) + " = " + oneline(translate_expr(m.group(2))))
            else:
                out.append(pad + oneline(translate_expr(st)))
        elif kind == "block":
            out.append(pad + "if True:")
            emit(node[1], ind + 1, out)
        elif kind == "if":
            _, cond, then, els = node
            c = cond.replace(" ", "")
            if c == "showdown":
                emit(then, ind, out)
            elif c == "!showdown":
                if els:
                    emit(els, ind, out)
            elif c == "log" or c == "pkmn.options.log":
                if els:
                    emit(els, ind, out)
            elif c == "!log":
                emit(then, ind, out)
            else:
                out.append(pad + "if " + oneline(translate_expr(cond)) + ":")
                emit(then, ind + 1, out)
                if els:
                    out.append(pad + "else:")
                    emit(els, ind + 1, out)
        elif kind == "loop":
            m2 = re.match(r"^for \(([\w.()]+), 0\.\.\) \|(\w+), (\w+)\|$", node[1])
            if m2:  # for (slice, 0..) |item, index|
                out.append(pad + f"for {m2.group(3)}, {m2.group(2)} in enumerate({oneline(translate_expr(m2.group(1)))}):")
                emit(node[2], ind + 1, out)
                continue
            m = re.match(r"^for \((.+)\.\.(.+)\) \|(\w+)\|$", node[1])
            if not m:
                raise Untranslatable("loop: " + node[1][:50])
            out.append(pad + f"for {m.group(3)} in range({oneline(translate_expr(m.group(1)))}, {oneline(translate_expr(m.group(2)))}):")
            emit(node[2], ind + 1, out)
    if len(out) == n0:
        out.append(pad + "pass")


def oneline(s):
    return re.sub(r"\s+", " ", s)


STATS_LOOP = re.compile(
    r"inline for \(@typeInfo\(@TypeOf\(t\.actual\.p1\.get\(1\)\.stats\)\)\.@\"struct\"\.fields\) \|f\| \{\s*"
    r"if \(!std\.mem\.eql\(u8, f\.name, \"hp\"\)\) \{\s*try expectEqual\(\s*@field\(([\w.()]+), f\.name\),\s*"
    r"@field\(([\w.()]+), f\.name\),\s*\);\s*\}\s*\}")


def main():
    src = open(REF).read()
    # the one comptime loop in the suite ("Transform effect"): every stat but hp of two Stats structs is equal
    src = STATS_LOOP.sub(lambda m: f"try expectStatsEqualExceptHp({m.group(1)}, {m.group(2)});", src)
    cases = []
    for m in re.finditer(r'^test "([^"]+)" \{', src, re.M):
        name = m.group(1)
        b0 = m.end() - 1
        body_src = strip_comments(src[b0:])
        be = match_close(body_src, 0, "{", "}")
        line = src.count("\n", 0, m.start()) + 1
        cases.append((name, line, body_src[1:be]))

    out = ['"""GENERATED by tools/transcribe_kat.py from the reference\'s gen1/test.zig — do not edit.',
           "",
           "Known-answer cases for the Generation I engine in Pokémon Showdown compatibility mode with",
           "libpkmn's default options (miss, advance, ebc), scripted roll lists. See tests/kat_harness.py.",
           '"""',
           "from kat_harness import *  # noqa: F401,F403",
           "", "CASES = []", ""]
    ok = bad = 0
    for idx, (name, line, body) in enumerate(cases):
        fn = f"case_{idx:03d}"
        try:
            nodes = parse_block(body, 0, len(body))
            lines = []
            emit(nodes, 1, lines)
            code = "\n".join(lines)
            compile("def f():\n" + code, fn, "exec")
            out.append(f"def {fn}():")
            out.append(f'    """{name} (gen1/test.zig:{line})"""')
            out.append(code)
            out.append("")
            out.append(f"CASES.append(({name!r}, {line}, {fn}, None))")
            ok += 1
        except (Untranslatable, SyntaxError, ValueError) as e:
            why = f"{type(e).__name__}: {e}"
            if os.environ.get("KAT_DEBUG") and isinstance(e, SyntaxError):
                print(name, e, "\n", "\n".join(lines[max(0, e.lineno - 3):e.lineno]))
            out.append(f"CASES.append(({name!r}, {line}, None, {why[:150]!r}))")
            bad += 1
        out.append("")
    path = os.path.join(ROOT, "tests", "golden", "gen1_kat_cases.py")
    open(path, "w").write("\n".join(out))
    print(f"transcribed {ok} of {len(cases)} tests ({bad} untranslatable) -> {path}")


if __name__ == "__main__":
    sys.exit(main())
