#!/usr/bin/env python
"""oracle/f2c_ref.py -- TEST INFRASTRUCTURE: builds oracle/_ref/libbmgref.so, the
REFERENCE's own serial BoxMG kernels, from the Fortran sources where they lie
under /root/reference/src/{2d,3d}/ftn (never copied into this repo).

The image has no Fortran compiler, so this recipe transliterates the small
Fortran subset those files use (block DO / IF, assignments, CALLs of LAPACK,
PARAMETER constants, explicit-shape arrays) statement by statement into C that
keeps the loop order and the expression order, writes the C into a temporary
directory, compiles it with `gcc -O2 -ffp-contract=off` and keeps ONLY the
shared library (oracle/_ref/ is git-ignored).  LAPACK (DPTTRF/DPTTRS/DPBTRF/
DPBTRS), a system dependency of the reference, is bound to the restatements in
oracle/bmg2.c (themselves checked against scipy's LAPACK).

Usage: python oracle/f2c_ref.py [--ref /root/reference] [--keep-c DIR]
"""
import argparse
import os
import re
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))

FILES_2D = ["BMG2_SymStd_SETUP_recip", "BMG2_SymStd_relax_GS", "BMG2_SymStd_residual", "BMG2_SymStd_restrict",
            "BMG2_SymStd_interp_add", "BMG2_SymStd_SETUP_interp_OI", "BMG2_SymStd_SETUP_ITLI_ex",
            "BMG2_SymStd_SETUP_lines_x", "BMG2_SymStd_SETUP_lines_y", "BMG2_SymStd_relax_lines_x",
            "BMG2_SymStd_relax_lines_y", "BMG2_SymStd_SETUP_cg_LU", "BMG2_SymStd_SOLVE_cg"]
FILES_3D = ["BMG3_SymStd_SETUP_recip", "BMG3_SymStd_relax_GS", "BMG3_SymStd_residual", "BMG3_SymStd_restrict",
            "BMG3_SymStd_interp_add", "BMG3_SymStd_SETUP_interp_OI", "BMG3_SymStd_SETUP_ITLI07_ex",
            "BMG3_SymStd_SETUP_ITLI27_ex", "BMG3_SymStd_SETUP_cg_LU", "BMG3_SymStd_SOLVE_cg"]

PRELUDE = r"""
#include <math.h>
#include <float.h>
#include <stddef.h>
#include <stdio.h>
#include <stdlib.h>
static inline int f_iabs(int a) { return a < 0 ? -a : a; }
#define f_abs(x) _Generic((x), double: fabs((double)(x)), float: fabs((double)(x)), default: f_iabs((int)(x)))
#define f_min2(a, b) ({ __typeof__((a) + (b)) a_ = (a), b_ = (b); a_ < b_ ? a_ : b_; })
#define f_max2(a, b) ({ __typeof__((a) + (b)) a_ = (a), b_ = (b); a_ > b_ ? a_ : b_; })
#define f_mod(a, b) ((a) % (b))
void dpttrf_(); void dpttrs_(); void dpbtrf_(); void dpbtrs_(); void dpotrf_(); void dpotrs_();
void print_error(const char *s);
"""

SHIM = r"""
#include <stdio.h>
#include <stdlib.h>
#include "oracle.h"
/* LAPACK entry points the Fortran calls, bound to the checked restatements */
void dpttrf_(int *n, double *d, double *e, int *info) { *info = orc_dpttrf(*n, d, e); }
void dpttrs_(int *n, int *nrhs, double *d, double *e, double *b, int *ldb, int *info)
{ for (int r = 0; r < *nrhs; r++) orc_dptts2(*n, d, e, b + (size_t)r * *ldb); *info = 0; }
void dpbtrf_(char *uplo, int *n, int *kd, double *ab, int *ldab, int *info)
{ (void)uplo; *info = orc_dpbtf2_upper(*n, *kd, ab, *ldab); }
void dpbtrs_(char *uplo, int *n, int *kd, int *nrhs, double *ab, int *ldab, double *b, int *ldb, int *info)
{ (void)uplo; for (int r = 0; r < *nrhs; r++) orc_dpbtrs_upper(*n, *kd, ab, *ldab, b + (size_t)r * *ldb); *info = 0; }
void dpotrf_(char *uplo, int *n, double *a, int *lda, int *info) { (void)uplo; *info = orc_dpotf2_upper(*n, a, *lda); }
void dpotrs_(char *uplo, int *n, int *nrhs, double *a, int *lda, double *b, int *ldb, int *info)
{ (void)uplo; for (int r = 0; r < *nrhs; r++) orc_dpotrs_upper(*n, a, *lda, b + (size_t)r * *ldb); *info = 0; }
void print_error(const char *s) { fprintf(stderr, "bmgref: %s\n", s); }
"""


def strip_and_join(text):
    """comments out, continuation lines joined, lower-cased outside strings"""
    out, cur = [], ""
    for raw in text.splitlines():
        line, instr, q = "", False, ""
        for ch in raw:
            if instr:
                line += ch
                if ch == q:
                    instr = False
            elif ch in "'\"":
                instr, q = True, ch
                line += ch
            elif ch == "!":
                break
            else:
                line += ch.lower()
        line = line.rstrip()
        if not line.strip():
            continue
        s = line.strip()
        if s.startswith("&"):
            s = s[1:].lstrip()
        cont = s.endswith("&")
        if cont:
            s = s[:-1].rstrip()
        cur += (" " if cur else "") + s
        if not cont:
            out.append(cur)
            cur = ""
    if cur:
        out.append(cur)
    return out


def split_top(s, sep=","):
    parts, depth, cur, instr = [], 0, "", False
    for ch in s:
        if ch in "'\"":
            instr = not instr
        if not instr:
            if ch == "(":
                depth += 1
            elif ch == ")":
                depth -= 1
            elif ch == sep and depth == 0:
                parts.append(cur.strip())
                cur = ""
                continue
        cur += ch
    if cur.strip():
        parts.append(cur.strip())
    return parts


def match_paren(s, i):
    """index of the parenthesis closing s[i] == '('"""
    depth = 0
    for k in range(i, len(s)):
        if s[k] == "(":
            depth += 1
        elif s[k] == ")":
            depth -= 1
            if depth == 0:
                return k
    raise ValueError("unbalanced: " + s)


def expr(e):
    """Fortran expression -> C (same operator order; intrinsics mapped)"""
    e = re.sub(r"(\d+\.?\d*|\.\d+)d([+-]?\d+)", r"\1e\2", e)           # 1.d0 -> 1.e0
    for f, c in ((".eq.", "=="), (".ne.", "!="), (".lt.", "<"), (".le.", "<="), (".gt.", ">"), (".ge.", ">="),
                 (".and.", "&&"), (".or.", "||"), (".not.", "!"), (".true.", "1"), (".false.", "0"), ("/=", "!=")):
        e = e.replace(f, " " + c + " ")
    if "**" in e:
        raise ValueError("power operator not supported: " + e)
    # n-ary min / max -> nested binary
    for name in ("min", "max"):
        while True:
            m = re.search(r"(?<![a-z0-9_])" + name + r"\s*\(", e)
            if not m:
                break
            close = match_paren(e, m.end() - 1)
            args = [a for a in split_top(e[m.end():close])]
            acc = args[0]
            for a in args[1:]:
                acc = f"f_{name}2({acc}, {a})"
            if len(args) == 1:
                acc = f"({acc})"
            e = e[:m.start()] + acc.replace("f_" + name + "2", "F_" + name.upper() + "2") + e[close + 1:]
        e = e.replace("F_" + name.upper() + "2", "f_" + name + "2")
    e = re.sub(r"(?<![a-z0-9_])iabs\s*\(", "f_iabs(", e)
    e = re.sub(r"(?<![a-z0-9_])d?abs\s*\(", "f_abs(", e)
    e = re.sub(r"(?<![a-z0-9_])mod\s*\(", "f_mod(", e)
    e = re.sub(r"(?<![a-z0-9_])epsilon\s*\([^)]*\)", "DBL_EPSILON", e)
    e = re.sub(r"(?<![a-z0-9_])(dble|real)\s*\(", "(double)(", e)
    return e


DECL_RE = re.compile(r"^(integer|real|double precision|logical|character)\s*(\*\s*\d+|\([^)]*\))?\s*(.*)$")


class Unit:
    """one SUBROUTINE -> one C function"""

    def __init__(self, lines, incdir):
        self.lines, self.incdir = lines, incdir
        self.consts = []          # (name, ctype, value) from PARAMETER statements
        self.types = {}           # name -> 'int' | 'double'
        self.dims = {}            # name -> [dim exprs]
        self.value_args = set()
        self.args, self.cname = [], None
        self.body = []
        self.indent = 1

    # ---- declarations -------------------------------------------------------
    def declare(self, base, rest):
        ctype = "int" if base in ("integer", "logical") else ("char" if base == "character" else "double")
        attrs, ents = "", rest
        if "::" in rest:
            attrs, ents = rest.split("::", 1)
        value = "value" in [a.strip() for a in split_top(attrs)] if attrs else False
        for ent in split_top(ents):
            m = re.match(r"^([a-z_][a-z0-9_]*)\s*(\((.*)\))?\s*(=\s*(.*))?$", ent)
            if not m:
                raise ValueError("declaration: " + ent)
            name = m.group(1)
            if name == "epsilon":
                continue          # typed intrinsic, not a variable
            self.types[name] = ctype
            if m.group(3) is not None:
                self.dims[name] = split_top(m.group(3))
            if value:
                self.value_args.add(name)
            if m.group(5) is not None:
                self.consts.append((name, ctype, m.group(5)))

    def parameter(self, inner):
        for item in split_top(inner):
            name, val = [t.strip() for t in item.split("=", 1)]
            self.consts.append((name, self.types.get(name, "int"), val))

    def include(self, fname):
        path = os.path.join(self.incdir, fname)
        for st in strip_and_join(open(path).read()):
            self.statement(st, in_include=True)

    # ---- statements ---------------------------------------------------------
    def emit(self, s):
        self.body.append("    " * self.indent + s)

    def call(self, st):
        m = re.match(r"^call\s+([a-z_0-9]+)\s*\((.*)\)$", st)
        name, args = m.group(1), split_top(m.group(2))
        cargs = []
        for a in args:
            if a[0] in "'\"":
                cargs.append('"' + a[1:-1].upper() + '"')
            elif re.match(r"^[a-z_][a-z0-9_]*$", a):
                cargs.append((a + "_") if a in self.dims else "&" + a)
            elif re.match(r"^[a-z_][a-z0-9_]*\s*\(", a) and a.split("(")[0].strip() in self.dims and \
                    match_paren(a, a.index("(")) == len(a) - 1:
                cargs.append("&" + expr(a))
            else:
                isreal = bool(re.search(r"\d\.|\.\d|[0-9]d[+-]?\d|(?<![a-z0-9_])r[a-z]+", a))
                cargs.append(f"&({'double' if isreal else 'int'}){{{expr(a)}}}")
        cname = name if name == "print_error" else name + "_"
        if name == "print_error":
            cargs = [c.lstrip("&") if not c.startswith('"') else c for c in cargs]
            cargs = ['"error reported by the reference kernel"']
        self.emit(f"{cname}({', '.join(cargs)});")

    def statement(self, st, in_include=False):
        m = re.match(r"^subroutine\s+([a-z_0-9]+)\s*\((.*?)\)\s*(bind\s*\(\s*c\s*,\s*name\s*=\s*['\"](.*?)['\"]\s*\))?$", st)
        if m:
            self.args = [a.strip() for a in split_top(m.group(2))]
            self.cname = m.group(4) or (m.group(1) + "_")
            return
        if re.match(r"^(use\s|implicit\s|external\s|end$|end\s+subroutine|contains)", st) or st == "end":
            return
        m = re.match(r"^include\s+['\"](.*)['\"]$", st)
        if m:
            return self.include(m.group(1))
        m = re.match(r"^parameter\s*\((.*)\)$", st)
        if m:
            return self.parameter(m.group(1))
        m = DECL_RE.match(st)
        if m and not re.match(r"^(real|integer)\s*\([^)]*\)\s*=", st):
            return self.declare(m.group(1), m.group(3))
        if in_include:
            raise ValueError("unexpected statement in include: " + st)
        # executable statements
        m = re.match(r"^do\s+([a-z_][a-z0-9_]*)\s*=\s*(.*)$", st)
        if m:
            v, parts = m.group(1), split_top(m.group(2))
            lo, hi = expr(parts[0]), expr(parts[1])
            step = expr(parts[2]) if len(parts) > 2 else "1"
            self.emit(f"{{ const int hi_ = ({hi}), st_ = ({step}); "
                      f"for ({v} = ({lo}); st_ > 0 ? {v} <= hi_ : {v} >= hi_; {v} += st_) {{")
            self.indent += 1
            return
        if re.match(r"^end\s*do$", st):
            self.indent -= 1
            return self.emit("} }")
        m = re.match(r"^(else\s*)?if\s*\(", st)
        if m:
            p0 = st.index("(", m.start())
            p1 = match_paren(st, p0)
            cond, tail = expr(st[p0 + 1:p1]), st[p1 + 1:].strip()
            if tail == "then":
                if m.group(1):
                    self.indent -= 1
                    self.emit(f"}} else if ({cond}) {{")
                else:
                    self.emit(f"if ({cond}) {{")
                self.indent += 1
            else:
                self.emit(f"if ({cond}) {{")
                self.indent += 1
                self.statement(tail)
                self.indent -= 1
                self.emit("}")
            return
        if st == "else":
            self.indent -= 1
            self.emit("} else {")
            self.indent += 1
            return
        if re.match(r"^end\s*if$", st):
            self.indent -= 1
            return self.emit("}")
        if st.startswith("call "):
            return self.call(st)
        if st == "return":
            return self.emit("return;")
        if st.startswith("write") or st.startswith("print") or re.match(r"^\d+\s+format", st):
            return self.emit("/* i/o statement dropped */;")
        if st == "stop":
            return self.emit("abort();")
        # assignment: split at the first top-level '='
        depth = 0
        for k, ch in enumerate(st):
            if ch == "(":
                depth += 1
            elif ch == ")":
                depth -= 1
            elif ch == "=" and depth == 0 and st[k + 1:k + 2] != "=" and st[k - 1] not in "<>/=":
                return self.emit(f"{expr(st[:k].strip())} = {expr(st[k + 1:].strip())};")
        raise ValueError("cannot translate: " + st)

    # ---- output -------------------------------------------------------------
    def render(self):
        for st in self.lines:
            self.statement(st)
        out = []
        seen = set()
        for name, ctype, val in self.consts:
            if name in seen:
                continue
            seen.add(name)
            out.append(f"static const {ctype} {name} = {expr(val)};")
        params, pre, post = [], [], []
        for a in self.args:
            t = self.types[a]
            if a in self.dims:
                params.append(f"{t} *{a}_")
            elif a in self.value_args:
                params.append(f"{t} {a}")
            else:
                params.append(f"{t} *{a}_p")
                pre.append(f"#define {a} (*{a}_p)")
                post.append(f"#undef {a}")
        out.append(f"void {self.cname}({', '.join(params)})\n{{")
        out += pre
        const_names = {c[0] for c in self.consts}
        for name, t in self.types.items():
            if name in self.args or name in const_names or t == "char":
                continue
            if name in self.dims:
                n = " * ".join(f"({expr(d)})" for d in self.dims[name])
                out.append(f"    {t} *{name}_ = ({t} *)calloc((size_t)({n}), sizeof({t}));")
                post.append(f"    free({name}_);")
            else:
                out.append(f"    {t} {name} = 0; (void){name};")
        for name, dims in self.dims.items():
            idx = [f"i{k}" for k in range(len(dims))]
            terms, stride = [], "1"
            for k, d in enumerate(dims):
                lo = "1"
                if ":" in d:
                    lo, d = [x.strip() for x in d.split(":")]
                terms.append(f"((size_t)(({idx[k]}) - ({expr(lo)}))) * {name}_s{k}")
                out.append(f"    const size_t {name}_s{k} = {stride}; (void){name}_s{k};")
                if d != "*":
                    stride = f"{name}_s{k} * (size_t)(({expr(d)}) - ({expr(lo)}) + 1)"
            out.append(f"#define {name}({', '.join(idx)}) {name}_[{' + '.join(terms)}]")
            post.append(f"#undef {name}")
        # `return` inside the body must still run the epilogue (frees): wrap body in a do-while
        body = "\n".join(self.body)
        if "free(" in "\n".join(post):
            body = body.replace("return;", "goto epilogue_;")
        out.append(body)
        out.append("epilogue_: ;")
        out += [p for p in post if p.startswith("    free")]
        out.append("}")
        out += [p for p in post if p.startswith("#undef")]
        return "\n".join(out) + "\n"


def translate(path):
    lines = strip_and_join(open(path).read())
    return Unit(lines, os.path.dirname(path)).render()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ref", default="/root/reference")
    ap.add_argument("--keep-c", default=None)
    a = ap.parse_args()
    if not os.path.isdir(os.path.join(a.ref, "src", "2d", "ftn")):
        print("f2c_ref: no reference tree at", a.ref, "-- nothing built")
        return 0
    outdir = os.path.join(HERE, "_ref")
    os.makedirs(outdir, exist_ok=True)
    tmp = a.keep_c or tempfile.mkdtemp(prefix="bmgref_")
    os.makedirs(tmp, exist_ok=True)
    srcs = []
    for sub, files in (("2d", FILES_2D), ("3d", FILES_3D)):
        for f in files:
            c = PRELUDE + translate(os.path.join(a.ref, "src", sub, "ftn", f + ".f90"))
            p = os.path.join(tmp, f + ".c")
            open(p, "w").write(c)
            srcs.append(p)
    shim = os.path.join(tmp, "lapack_shim.c")
    open(shim, "w").write(SHIM)
    cmd = ["gcc", "-O2", "-march=x86-64-v3", "-ffp-contract=off", "-fno-fast-math", "-fPIC", "-shared", "-w",
           "-I", HERE, "-o", os.path.join(outdir, "libbmgref.so")] + srcs + [shim, os.path.join(HERE, "bmg2.c"), "-lm",
                                                                            "-fopenmp"]
    subprocess.check_call(cmd)
    if not a.keep_c:
        for p in srcs + [shim]:
            os.remove(p)
        os.rmdir(tmp)
    print("f2c_ref: built", os.path.join(outdir, "libbmgref.so"), "from", len(srcs), "reference kernels")
    return 0


if __name__ == "__main__":
    sys.exit(main())
