#!/usr/bin/env python
"""Build the API reference from the sources (the reference does this with doxygen + sphinx, doc/Doxyfile, doc/conf.py;
neither tool is needed here):

    python scripts/gen_api_doc.py            # writes doc/cpp_api.md, doc/c_api.md, doc/python_api.md, doc/index.md
    python scripts/gen_api_doc.py --check    # exit 1 if the committed files are not what the sources produce

C/C++: every declaration of the public headers with the comment in front of it (`/*! ... */`, `/* ... */` or `//`),
grouped by the section rulers the headers already carry. Python: signature + docstring of every public name.
"""
import ast
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CPP_HEADERS = ["include/rabit/rabit.h", "include/rabit/serializable.h", "include/rabit/internal/engine.h",
               "include/rabit/internal/io.h", "include/rabit/internal/utils.h", "include/rabit/internal/timer.h"]
C_HEADERS = ["include/rabit/c_api.h", "include/rabit_device.h"]
PY_MODULES = ["rabit_b200/rabit.py", "rabit_b200/torch_binding.py", "rabit_b200/parallel/launcher.py",
              "rabit_b200/parallel/tracker.py", "rabit_b200/models/histogram.py", "rabit_b200/models/checkpoint.py",
              "rabit_b200/ops/emulate.py", "rabit_b200/utils/timing.py"]

RULER = re.compile(r"^\s*(?://|/\*)\s*-{2,}\s*(.*?)\s*-*\s*(?:\*/)?\s*$")


def clean_comment(lines):
    """comment lines -> markdown paragraph(s); doxygen commands become bold labels / bullets"""
    text = []
    for ln in lines:
        ln = ln.strip()
        ln = re.sub(r"^/\*[!*]?\s?", "", ln)
        ln = re.sub(r"\s*\*/$", "", ln)
        ln = re.sub(r"^\*\s?", "", ln)
        ln = re.sub(r"^//[/!]?\s?", "", ln)
        text.append(ln)
    out, bullets = [], False
    for ln in text:
        m = re.match(r"[\\@](t?param)\s+(\S+)\s*(.*)", ln)
        if m:
            out.append("* `%s` — %s" % (m.group(2), m.group(3)))
            bullets = True
            continue
        m = re.match(r"[\\@](return|returns)\s*(.*)", ln)
        if m:
            out.append("* **returns** " + m.group(2))
            bullets = True
            continue
        m = re.match(r"[\\@](sa|note)\s*(.*)", ln)
        if m:
            out.append("* **%s** %s" % ("see also" if m.group(1) == "sa" else "note", m.group(2)))
            bullets = True
            continue
        ln = re.sub(r"^[\\@]brief\s*", "", ln)
        if bullets and ln and out:
            out[-1] += " " + ln.strip()     # continuation of a bullet
        else:
            out.append(ln)
            bullets = False
    return "\n".join(out).strip()


def parse_header(path):
    """-> (intro, [(section, [(decl, comment)])])"""
    lines = open(os.path.join(ROOT, path)).read().split("\n")
    i, n = 0, len(lines)
    intro = None
    sections = [("", [])]
    pending = []         # comment lines waiting for their declaration
    depth_skip = 0       # inside a body we do not document ({ ... } of inline definitions)
    hidden = False       # private / protected part of a class
    while i < n:
        ln = lines[i]
        s = ln.strip()
        if depth_skip:
            depth_skip += s.count("{") - s.count("}")
            i += 1
            continue
        m = RULER.match(ln)
        if m and m.group(1):
            sections.append((m.group(1), []))
            pending = []
            i += 1
            continue
        if s.startswith("/*"):
            block = [ln]
            while "*/" not in lines[i]:
                i += 1
                block.append(lines[i])
            i += 1
            if intro is None and not any(e for _, e in sections):
                intro = clean_comment(block)
                pending = []
            else:
                pending = block
            continue
        if s.startswith("//"):
            pending = pending + [ln] if pending and pending[-1].strip().startswith("//") else [ln]
            i += 1
            continue
        if not s:
            pending = []
            i += 1
            continue
        if s.startswith("#") or s.startswith("namespace") or s.startswith("}") or s.startswith("extern \"C\"") or \
                s.startswith("using ") or s in ("public:", "private:", "protected:"):
            if s.startswith("#define") and s.endswith("\\"):
                while lines[i].rstrip().endswith("\\"):
                    i += 1
            i += 1
            if s in ("private:", "protected:"):
                hidden = True
            elif s == "public:" or s.startswith("}"):
                hidden = False
            pending = []
            continue
        # a declaration: up to the first ';' or '{' at paren depth 0
        decl, paren = [], 0
        body = False
        while i < n:
            cur = lines[i]
            code = cur.split("//")[0] if "//" in cur and "://" not in cur else cur
            code = re.sub(r"/\*.*?\*/", "", code)
            decl.append(code.rstrip())
            paren += code.count("(") - code.count(")")
            t = code.strip()
            i += 1
            if paren <= 0 and (t.endswith(";") or t.endswith("{") or t.endswith("}")):
                body = t.endswith("{")
                opened = sum(x.count("{") - x.count("}") for x in decl)
                if opened > 0:
                    first = "\n".join(decl)
                    if re.match(r"\s*(template\s*<[^>]*>\s*)?(class|struct|enum)\b", first):
                        body = False          # walk INTO classes: their members are API
                    else:
                        depth_skip = opened
                break
        text = "\n".join(decl).strip()
        text = re.sub(r"\s*\{\s*$", "", text) if (body or text.endswith("{")) else text
        text = re.sub(r",?\s*const char \*_file = _FILE,\s*const int _line = _LINE,\s*const char \*_caller = _CALLER", "", text)
        if "{" in text and text.rstrip().endswith("}") and not re.match(r"\s*(typedef\s+)?enum\b", text):
            text = text[:text.index("{")].rstrip()      # one-line inline definition: keep the signature
        if text and not text.startswith("}") and not hidden:
            sections[-1][1].append((text, clean_comment(pending) if pending else ""))
        pending = []
    return intro or "", [(t, e) for t, e in sections if e]


def header_markdown(paths, title, lead):
    out = ["# " + title, "", lead, ""]
    for p in paths:
        intro, sections = parse_header(p)
        out += ["## `%s`" % p, "", intro, ""]
        for name, entries in sections:
            if name:
                out += ["### " + name, ""]
            for decl, comment in entries:
                out += ["```cpp", decl, "```"]
                if comment:
                    out += [comment]
                out += [""]
    return "\n".join(out).rstrip() + "\n"


def py_signature(fn):
    a = fn.args
    parts = []
    pos = a.posonlyargs + a.args
    defaults = [None] * (len(pos) - len(a.defaults)) + list(a.defaults)
    for arg, d in zip(pos, defaults):
        parts.append(arg.arg + ("=" + ast.unparse(d) if d is not None else ""))
    if a.vararg:
        parts.append("*" + a.vararg.arg)
    for arg, d in zip(a.kwonlyargs, a.kw_defaults):
        parts.append(arg.arg + ("=" + ast.unparse(d) if d is not None else ""))
    if a.kwarg:
        parts.append("**" + a.kwarg.arg)
    return "%s(%s)" % (fn.name, ", ".join(parts))


def python_markdown():
    out = ["# Python API", "", "Generated from the docstrings (`scripts/gen_api_doc.py`). `rabit_b200.rabit` is call-compatible with the "
           "reference's `python/rabit.py`; everything else is new.", ""]
    for p in PY_MODULES:
        tree = ast.parse(open(os.path.join(ROOT, p)).read())
        mod = p[:-3].replace("/", ".")
        out += ["## `%s`" % mod, "", (ast.get_docstring(tree) or "").strip(), ""]
        consts = []
        for node in tree.body:
            if isinstance(node, ast.Assign) and len(node.targets) == 1 and isinstance(node.targets[0], ast.Name):
                name = node.targets[0].id
                if name.isupper() and not name.startswith("_") and isinstance(node.value, ast.Constant):
                    consts.append("`%s = %r`" % (name, node.value.value))
        if consts:
            out += ["Constants: " + ", ".join(consts), ""]
        for node in tree.body:
            if isinstance(node, ast.FunctionDef) and not node.name.startswith("_"):
                out += ["### `%s`" % py_signature(node), "", (ast.get_docstring(node) or "").strip(), ""]
            elif isinstance(node, ast.ClassDef) and not node.name.startswith("_"):
                out += ["### class `%s`" % node.name, "", (ast.get_docstring(node) or "").strip(), ""]
                for sub in node.body:
                    if isinstance(sub, ast.FunctionDef) and (not sub.name.startswith("_") or sub.name == "__init__"):
                        doc = (ast.get_docstring(sub) or "").strip()
                        out += ["* `%s`%s" % (py_signature(sub), " — " + " ".join(doc.split()) if doc else "")]
                out += [""]
    return "\n".join(out).rstrip() + "\n"


INDEX = """# rabit-b200 documentation

* [Tutorial](guide.md) — the programming model, fault tolerance walk-through, lazy allreduce, local models, device buffers
* [Parameters](parameters.md) — every `name=value` parameter and environment variable
* [API overview](api.md) — the public surface at a glance, with what is new next to the reference
* API reference, generated from the sources by `python scripts/gen_api_doc.py`:
  [C++](cpp_api.md) · [C](c_api.md) · [Python](python_api.md)
* [`DESIGN.md`](../DESIGN.md) — architecture and measurements · [`INTEGRATION.md`](../INTEGRATION.md) — embedding, limits
"""


def main():
    files = {
        "doc/cpp_api.md": header_markdown(
            CPP_HEADERS, "C++ API reference",
            "Generated from the public headers (`scripts/gen_api_doc.py`). The trailing `_file, _line, _caller` arguments "
            "that every collective takes (defaulted to the call site, used in debug logs and cache keys) are left out."),
        "doc/c_api.md": header_markdown(
            C_HEADERS, "C API reference", "Generated from the public headers (`scripts/gen_api_doc.py`)."),
        "doc/python_api.md": python_markdown(),
        "doc/index.md": INDEX,
    }
    check = "--check" in sys.argv
    stale = []
    for path, text in files.items():
        full = os.path.join(ROOT, path)
        if check:
            if not os.path.exists(full) or open(full).read() != text:
                stale.append(path)
        else:
            with open(full, "w") as f:
                f.write(text)
            print("wrote %s (%d lines)" % (path, text.count("\n")))
    if stale:
        print("stale: %s (run python scripts/gen_api_doc.py)" % ", ".join(stale))
        sys.exit(1)


if __name__ == "__main__":
    main()
