"""Compare two builds of libfa_b200.so kernel by kernel: python tools/sass_diff.py old.so new.so
Kernels are matched by name + template arguments (namespaces ignored); instruction text and encodings must be identical
(whitespace of the listing ignored).  Used to show that a refactor or a new instantiation left the measured kernels untouched."""
import re
import subprocess
import sys


def kernels(lib):
    out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
    res, name, buf = {}, None, []
    for line in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            if name:
                res[name] = buf
            name, buf = m.group(1), []
        elif name and re.match(r"\s*/\*", line):
            buf.append(re.sub(r"\s+", " ", line.strip()))
    if name:
        res[name] = buf
    return res


def keys(mangled_names):
    """kernel name + template arguments; the enclosing namespaces (an anonymous one carries a per-object hash) are dropped"""
    names = list(mangled_names)
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True, check=True).stdout.splitlines()
    res = {}
    for m, d in zip(names, out):
        hit = re.search(r"(\w+)(<[^()]*>)?\(", d.replace("(anonymous namespace)::", ""))
        res[m] = (hit.group(1) + (hit.group(2) or "")) if hit else m
    return res


def main():
    a, b = kernels(sys.argv[1]), kernels(sys.argv[2])
    ka, kb = keys(a), keys(b)
    a, b = {ka[k]: v for k, v in a.items()}, {kb[k]: v for k, v in b.items()}
    gone, new = sorted(set(a) - set(b)), sorted(set(b) - set(a))
    changed = sorted(k for k in a if k in b and a[k] != b[k])
    print(f"{len(a)} kernels in {sys.argv[1]}, {len(b)} in {sys.argv[2]}: {len(set(a) & set(b)) - len(changed)} identical, "
          f"{len(changed)} changed, {len(gone)} gone, {len(new)} new")
    for tag, names in (("changed", changed), ("gone", gone), ("new", new)):
        for n in names[:40]:
            print(f"  {tag}: {n}")
    return 1 if changed or gone else 0


if __name__ == "__main__":
    sys.exit(main())
