"""integration/NativeMethods.cs is the P/Invoke binding a maintainer adds to the reference's C# project.
It cannot be compiled here (no .NET toolchain), so this test parses it: every [StructLayout(Sequential)] struct must list the
fields of the matching ctypes struct (storage_b200/_abi.py, which the GPU tests run on) in the same order with types of the
same size, which gives the same total size under the common alignment rules; every function include/lsmc_b200.h declares must
have a DllImport with the same number of parameters."""
import ctypes as C
import os
import re

from storage_b200 import _abi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CS = open(os.path.join(ROOT, "integration", "NativeMethods.cs")).read()
HEADER = re.sub(r"/\*.*?\*/", "", open(os.path.join(ROOT, "include", "lsmc_b200.h")).read(), flags=re.S)

CS_SIZES = {"int": 4, "uint": 4, "long": 8, "ulong": 8, "double": 8, "IntPtr": 8}
STRUCTS = {"LsmcStorage": _abi.lsmc_storage, "LsmcProblem": _abi.lsmc_problem, "LsmcMultiFactorParams": _abi.lsmc_multi_factor_params,
           "LsmcTriggerPrices": _abi.lsmc_trigger_prices, "LsmcResults": _abi.lsmc_results}


def cs_struct_fields(name):
    body = re.search(r"internal struct %s\s*\{(.*?)\n        \}" % name, CS, flags=re.S).group(1)
    fields = []
    for decl in re.findall(r"public\s+(fixed\s+)?([A-Za-z]+\*?)\s+(\w+)(?:\[(\d+)\])?;", body):
        fixed, ctype, fname, count = decl
        if ctype in STRUCTS:
            size, align = C.sizeof(STRUCTS[ctype]), C.alignment(STRUCTS[ctype])
        elif ctype.endswith("*"):
            size = align = 8
        else:
            size = align = CS_SIZES[ctype]
        fields.append((fname, size * (int(count) if count else 1), align))
    return fields


def layout_size(fields):
    offset, max_align = 0, 1
    for _, size, align in fields:
        offset = (offset + align - 1) // align * align + size
        max_align = max(max_align, align)
    return (offset + max_align - 1) // max_align * max_align


def snake(name):
    return re.sub(r"(?<!^)(?=[A-Z])", "_", name).lower()


def test_struct_layouts_match_the_ctypes_structs():
    for cs_name, ct in STRUCTS.items():
        fields = cs_struct_fields(cs_name)
        assert [snake(f[0]) for f in fields] == [n for n, _ in ct._fields_], cs_name
        for (fname, size, _), (cname, ctype) in zip(fields, ct._fields_):
            assert size == C.sizeof(ctype), f"{cs_name}.{fname}: {size} bytes in C#, {C.sizeof(ctype)} in the header"
        assert layout_size(fields) == C.sizeof(ct), cs_name


def header_functions():
    out = {}
    for m in re.finditer(r"\b(lsmc_[a-z_0-9]+)\s*\(([^;{]*?)\)\s*;", HEADER):
        name, params = m.group(1), m.group(2).strip()
        if name == "lsmc_progress_fn":
            continue
        out[name] = 0 if params in ("", "void") else params.count(",") + 1
    return out


def test_every_exported_function_has_a_dllimport_with_the_same_arity():
    declared = header_functions()
    assert set(declared) == {n for n, _, _ in _abi.SYMBOLS}
    imports = {}
    for m in re.finditer(r"\[DllImport\(Lib[^\]]*\)\]\s*internal static extern \S+ (lsmc_\w+)\(([^)]*)\);", CS):
        params = m.group(2).strip()
        imports[m.group(1)] = 0 if not params else params.count(",") + 1
    assert set(imports) == set(declared), set(imports) ^ set(declared)
    for name, arity in declared.items():
        assert imports[name] == arity, f"{name}: {imports[name]} parameters in C#, {arity} in the header"


def test_constants_match():
    assert "CommIdBytes = %d" % _abi.LSMC_COMM_ID_BYTES in CS
    assert "StreamPanels = 1u << 18" in CS and _abi.LSMC_STREAM_PANELS == 1 << 18
    for name, value in (("MaxBasis", 16), ("MaxFactors", 8), ("MaxDecisions", 16), ("NumTriggerVolumes", 10)):
        assert re.search(r"%s = %d\b" % (name, value), CS)
