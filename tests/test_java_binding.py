"""CPU: source-level checks of the Java half of the drop-in (radio-sim_b200/java/).  This image has no JDK, so the
adapter cannot be compiled here; these tests catch what a compiler would catch at the boundary with the reference:

 * every method called on a reference class (Simulator, Node, Transciever, RadioPacket, Position, the media, the
   event classes) exists in the reference SOURCE with that name and number of arguments
   (tests/golden/reference_java_api.json, regenerated and compared while /root/reference is present) -- round 1's
   adapter called RadioPacket.getTxpower(), which does not exist;
 * every struct offset constant in CudaRadioMedium.java equals offsetof()/sizeof() of include/radiomedium_b200.h
   under gcc;
 * every Panama downcall descriptor has the parameter count of the C prototype it binds;
 * java/reference-edits.patch applies to the reference tree, and what it calls on CudaRadioMedium is declared there.
"""
import json
import os
import re
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
JAVA = os.path.join(ROOT, "radio-sim_b200", "java")
ADAPTER = os.path.join(JAVA, "se", "sics", "emul8", "radiomedium", "CudaRadioMedium.java")
BULK = os.path.join(JAVA, "se", "sics", "emul8", "radiomedium", "events", "BulkReceptionEvent.java")
GENGOLDEN = os.path.join(JAVA, "tools", "GenGolden.java")
PATCH = os.path.join(JAVA, "reference-edits.patch")
HEADER = os.path.join(ROOT, "include", "radiomedium_b200.h")
API_JSON = os.path.join(ROOT, "tests", "golden", "reference_java_api.json")
REFERENCE = "/root/reference/radio-medium"


def strip_comments(text):
    text = re.sub(r"/\*.*?\*/", lambda m: " " * len(m.group(0)), text, flags=re.S)
    text = re.sub(r"//[^\n]*", "", text)
    return re.sub(r'"(?:\\.|[^"\\])*"', '""', text)


def api():
    a = json.load(open(API_JSON))
    # what the patch adds to the reference classes
    added = [l[1:] for l in open(PATCH).read().splitlines() if l.startswith("+") and not l.startswith("+++")]
    for ret, name, params in re.findall(r"public\s+([\w\[\]]+)\s+(\w+)\s*\(([^)]*)", "\n".join(added).replace("\n", " ")):
        a["Simulator"]["methods"].setdefault(name, []).append([ret, params.count(",") + 1 if params.strip() else 0])
    return a


def lookup(a, cls, name):
    """(return type, arities) of method `name` on class `cls`, following `extends`."""
    seen = set()
    while cls in a and cls not in seen:
        seen.add(cls)
        m = a[cls]["methods"].get(name)
        if m:
            return m[0][0], sorted(x[1] for x in m)
        cls = a[cls].get("extends")
    return None


def split_args(s):
    depth, n, cur = 0, 0, ""
    for ch in s:
        if ch in "([{":
            depth += 1
        elif ch in ")]}":
            depth -= 1
        if ch == "," and depth == 0:
            n += 1
            cur = ""
        else:
            cur += ch
    return n + (1 if cur.strip() else 0)


def match_back(text, i):
    """index of the '(' or '[' that matches the closer at text[i]"""
    close, opn = text[i], {")": "(", "]": "["}[text[i]]
    depth = 0
    while i >= 0:
        if text[i] == close:
            depth += 1
        elif text[i] == opn:
            depth -= 1
            if depth == 0:
                return i
        i -= 1
    raise ValueError("unbalanced")


def match_fwd(text, i):
    depth = 0
    for j in range(i, len(text)):
        if text[j] == "(":
            depth += 1
        elif text[j] == ")":
            depth -= 1
            if depth == 0:
                return j
    raise ValueError("unbalanced")


def declared_types(text, a):
    """variable / field / parameter name -> declared type, for the reference's class names only"""
    types = {}
    names = "|".join(sorted(a, key=len, reverse=True))
    for typ, arr, var in re.findall(r"\b(%s)\b(\s*\[\])?\s+(\w+)\s*(?=[=;,):])" % names, text):
        types[var] = typ + ("[]" if arr else "")
    types.setdefault("simulator", "Simulator")          # AbstractRadioMedium.java:37 (inherited field)
    return types


def receiver_type(text, dot, types, a, self_cls):
    """type of the expression that ends just before text[dot] == '.' (None = not a reference class)"""
    i = dot - 1
    while i >= 0 and text[i].isspace():
        i -= 1
    if text[i] == ")":
        o = match_back(text, i)
        j = o - 1
        while j >= 0 and text[j].isspace():
            j -= 1
        k = j
        while k >= 0 and (text[k].isalnum() or text[k] == "_"):
            k -= 1
        name = text[k + 1:j + 1]
        if not name:                                        # a parenthesised expression: look for a cast inside
            m = re.match(r"\(\s*\((\w+)\)", text[o:i + 1])
            return m.group(1) if m and m.group(1) in a else None
        p = k
        while p >= 0 and text[p].isspace():
            p -= 1
        if p >= 0 and text[p] == ".":
            t = receiver_type(text, p, types, a, self_cls)
        else:
            t = self_cls                                    # unqualified call: a method of this class or its parents
        if t is None:
            return None
        r = lookup(a, t.rstrip("[]"), name)
        return r[0] if r else None
    if text[i] == "]":
        o = match_back(text, i)
        j = o - 1
        k = j
        while k >= 0 and (text[k].isalnum() or text[k] == "_"):
            k -= 1
        t = types.get(text[k + 1:j + 1])
        return t[:-2] if t and t.endswith("[]") else None
    k = i
    while k >= 0 and (text[k].isalnum() or text[k] == "_"):
        k -= 1
    name = text[k + 1:i + 1]
    if name == "this":
        return self_cls
    return types.get(name)


def check_calls(path, self_cls, a):
    """-> list of (class, method, nargs) resolved on reference classes; asserts each exists"""
    text = strip_comments(open(path).read())
    types = declared_types(text, a)
    found = []
    # methods the file's own classes declare (fixture entries with constructors == [] and no methods are "own")
    own = {c for c in a if not a[c]["methods"] and not a[c]["fields"]}
    own_methods = set(re.findall(r"[\w<>\[\]]+\s+(\w+)\s*\([^()]*\)\s*(?:throws [\w, ]+)?\s*\{", text))
    for m in re.finditer(r"\.\s*(\w+)\s*\(", text):
        t = receiver_type(text, m.start(), types, a, self_cls)
        if t is None or t.endswith("[]") or t not in a:
            continue
        name = m.group(1)
        close = match_fwd(text, m.end() - 1)
        n = split_args(text[m.end():close])
        r = lookup(a, t, name)
        line = text.count("\n", 0, m.start()) + 1
        if r is None and t in own and name in own_methods:
            continue
        assert r is not None, f"{os.path.basename(path)}:{line}: {t}.{name}() is not declared in the reference"
        assert n in r[1], f"{os.path.basename(path)}:{line}: {t}.{name} takes {r[1]} arguments, called with {n}"
        found.append((t, name, n))
    # public fields read directly (Position.x / y / z)
    for m in re.finditer(r"\b(\w+)\.(\w+)\b(?!\s*\()", text):
        t = types.get(m.group(1))
        if t in a and not t.endswith("[]") and t not in own:
            assert m.group(2) in a[t]["fields"], f"{os.path.basename(path)}: {t}.{m.group(2)} is not a public field"
    return found


def test_reference_api_fixture_is_current():
    if not os.path.isdir(REFERENCE):
        pytest.skip("reference tree not present")
    import importlib.util
    spec = importlib.util.spec_from_file_location("mk", os.path.join(ROOT, "tests", "golden", "make_reference_java_api.py"))
    mk = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mk)
    assert json.loads(json.dumps(mk.build())) == json.load(open(API_JSON))


def test_the_resolver_rejects_a_wrong_method(tmp_path):
    bad = tmp_path / "Bad.java"
    bad.write_text("class Bad { void f(RadioPacket p, Node[] nodes) { double a = p.getTxpower(); } }")
    with pytest.raises(AssertionError, match="getTxpower"):
        check_calls(str(bad), "Bad", api())
    bad.write_text("class Bad { void f(Node[] nodes) { double a = nodes[0].getRadio().getTxpower(); } }")
    with pytest.raises(AssertionError, match="Transciever.getTxpower"):
        check_calls(str(bad), "Bad", api())
    bad.write_text("class Bad { void f(Simulator s, RadioPacket p, Node n) { s.generateReceptionEvents(p, n, 1.0); } }")
    with pytest.raises(AssertionError, match="arguments"):
        check_calls(str(bad), "Bad", api())


def test_adapter_calls_only_what_the_reference_declares():
    a = api()
    a["CudaRadioMedium"] = dict(methods={}, fields={}, constructors=[], extends="AbstractRadioMedium")
    found = check_calls(ADAPTER, "CudaRadioMedium", a)
    must = {("RadioPacket", "getTransmitPower", 0), ("RadioPacket", "getWirelessChannel", 0), ("RadioPacket", "getSource", 0),
            ("Transciever", "getTransmitPower", 0), ("Transciever", "getRxProbability", 0), ("Transciever", "getTxProbability", 0),
            ("Transciever", "isEnabled", 0), ("Node", "getPosition", 0), ("Node", "getRadio", 0), ("Node", "getIdAsInteger", 0),
            ("Simulator", "getNodes", 0), ("Simulator", "getRandom", 0), ("Simulator", "generateReceptionEventsBulk", 9),
            ("Simulator", "generateReceptionEvents", 4), ("Simulator", "generateTransmissionEvents", 1),
            ("Simulator", "deliverRadioPacket", 3)}
    assert must <= set(found), must - set(found)


def test_bulk_event_and_gengolden_call_only_what_the_reference_declares():
    a = api()
    a["BulkReceptionEvent"] = dict(methods={}, fields={}, constructors=[], extends="TimeEvent")
    found = check_calls(BULK, "BulkReceptionEvent", a)
    assert {("Transciever", "setReceiving", 2), ("Transciever", "clearReceiving", 0), ("Simulator", "deliverRadioPacket", 3)} <= set(found)
    a["GenGolden"] = dict(methods={}, fields={}, constructors=[], extends=None)
    a["RecordingSimulator"] = dict(methods={}, fields={}, constructors=[], extends="Simulator")
    found = check_calls(GENGOLDEN, "GenGolden", a)
    assert {("RecordingSimulator", "addNode", 2), ("Position", "set", 2), ("Transciever", "setRxProbability", 1),
            ("UDGMRadioMedium", "setSuccessRatioRx", 1), ("RadioMedium", "transmit", 1), ("RadioPacket", "setWirelessChannel", 1)} <= set(found)
    # constructors GenGolden uses
    assert 3 in a["RadioPacket"]["constructors"] and 1 in a["Simulator"]["constructors"] and 1 in a["N2NRadioMedium"]["constructors"]


def test_overrides_match_the_interface():
    """@Override methods of the adapter exist with the same arity in RadioMedium / AbstractRadioMedium / TimeEvent."""
    a = api()
    for path, parent in ((ADAPTER, "AbstractRadioMedium"), (BULK, "TimeEvent")):
        text = strip_comments(open(path).read())
        for ret, name, params in re.findall(r"@Override\s+public\s+(?:synchronized\s+)?(\w+)\s+(\w+)\s*\(([^)]*)\)", text):
            if name == "toString":
                continue
            r = lookup(a, parent, name) or lookup(a, "RadioMedium", name)
            assert r is not None, f"{name} overrides nothing"
            assert (params.count(",") + 1 if params.strip() else 0) in r[1] and r[0] == ret, name
    text = strip_comments(open(ADAPTER).read())
    declared = set(re.findall(r"public\s+(?:synchronized\s+)?\w+\s+(\w+)\s*\(", text))
    assert {"getName", "transmit", "getBaseRSSI"} <= declared          # RadioMedium.java:35-45 (setSimulator is inherited)


def test_struct_offsets_match_the_header(tmp_path):
    text = strip_comments(open(ADAPTER).read())
    consts = re.findall(r"\b(CFG|RES)_(\w+)\s*=\s*(\d+)", text)
    assert len(consts) >= 20
    struct = {"CFG": "rmb_config", "RES": "rmb_result"}
    lines = []
    for pre, name, _ in consts:
        if name == "SIZE":
            lines.append(f'printf("%zu\\n", sizeof({struct[pre]}));')
        else:
            lines.append(f'printf("%zu\\n", offsetof({struct[pre]}, {name.lower()}));')
    src = tmp_path / "off.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "radiomedium_b200.h"\nint main(void){\n' + "\n".join(lines) + "\nreturn 0;}\n")
    exe = tmp_path / "off"
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    got = [int(v) for v in subprocess.check_output([str(exe)], text=True).split()]
    for (pre, name, val), g in zip(consts, got):
        assert int(val) == g, f"{pre}_{name} = {val} in CudaRadioMedium.java, the header says {g}"
    # enum values used by the adapter
    hdr = open(HEADER).read()
    for cname, ename in (("MODEL_SINR", "RMB_MODEL_SINR"), ("MODEL_N2N", "RMB_MODEL_N2N"), ("MODEL_UDGM_CONSTANT", "RMB_MODEL_UDGM_CONSTANT"),
                         ("FLAG_COMPACT", "RMB_FLAG_COMPACT"), ("FLAG_F64_VALUES", "RMB_FLAG_F64_VALUES"),
                         ("RMB_ERR_UNIFORMS", "RMB_ERR_UNIFORMS"), ("DELIVERED", "RMB_DELIVERED")):
        j = int(re.search(r"\b%s\s*=\s*(-?\d+)" % cname, text).group(1))
        c = int(re.search(r"\b%s\s*=\s*(-?\d+)" % ename, hdr).group(1))
        assert j == c, cname


def test_downcall_descriptors_have_the_arity_of_the_prototypes():
    text = strip_comments(open(ADAPTER).read().replace('"rmb_', '"@rmb_'))       # keep the symbol names through the string stripping
    raw = open(ADAPTER).read()
    hdr = re.sub(r"/\*.*?\*/", "", open(HEADER).read(), flags=re.S)
    n = 0
    for m in re.finditer(r'fn\("(rmb_\w+)",\s*FunctionDescriptor\.(of|ofVoid)\(', raw):
        sym, kind = m.group(1), m.group(2)
        close = match_fwd(raw, m.end() - 1)
        layouts = [s.strip() for s in raw[m.end():close].split(",")]
        params = layouts[1:] if kind == "of" else layouts
        proto = re.search(r"\b%s\s*\(([^;]*?)\)\s*;" % sym, hdr, flags=re.S)
        assert proto, sym
        cparams = [p.strip() for p in proto.group(1).split(",")]
        assert len(params) == len(cparams), f"{sym}: {len(params)} layouts for {len(cparams)} C parameters"
        for lay, cp in zip(params, cparams):
            if "*" in cp:
                assert lay == "ADDRESS", (sym, cp, lay)
            elif "size_t" in cp or "int64_t" in cp:
                assert lay == "JAVA_LONG", (sym, cp, lay)
            elif "int32_t" in cp or cp.startswith("int "):
                assert lay == "JAVA_INT", (sym, cp, lay)
        n += 1
    assert n >= 10 and text


def test_java_sources_are_balanced():
    for path in (ADAPTER, BULK, GENGOLDEN):
        text = strip_comments(open(path).read())
        for o, c in ("{}", "()", "[]"):
            assert text.count(o) == text.count(c), (os.path.basename(path), o)


def test_reference_edits_patch_applies_and_uses_declared_adapter_methods(tmp_path):
    patch = open(PATCH).read()
    adapter = strip_comments(open(ADAPTER).read())
    declared = set(re.findall(r"public\s+(?:synchronized\s+)?[\w\[\]]+\s+(\w+)\s*\(", adapter)) | {"CudaRadioMedium"}
    added = "\n".join(l[1:] for l in patch.splitlines() if l.startswith("+") and not l.startswith("+++"))
    for name in re.findall(r"\(\(CudaRadioMedium\)[^)]*\)\s*\)?\.(\w+)\(", added) + re.findall(r"\bcuda\.(\w+)\(", added):
        assert name in declared, f"the patch calls CudaRadioMedium.{name}(), which the adapter does not declare"
    for const in re.findall(r"CudaRadioMedium\.(MODEL_\w+)", added):
        assert re.search(r"\b%s\s*=" % const, adapter), const
    # the hook is the FIRST statement of emulatorTimeStepDone, before `currentTime = stepTime`
    assert re.search(r"emulatorTimeStepDone\(\) \{\n(\+[^\n]*\n)+ +this\.currentTime = stepTime;", patch)
    if not os.path.isdir(REFERENCE) or not shutil.which("patch"):
        pytest.skip("reference tree (or patch) not present")
    work = tmp_path / "ref"
    shutil.copytree(os.path.join(REFERENCE, "java"), work / "radio-medium" / "java")
    subprocess.check_call(["patch", "-p1", "-s", "-i", PATCH], cwd=work)
    sim = open(work / "radio-medium" / "java" / "se" / "sics" / "emul8" / "radiomedium" / "Simulator.java").read()
    assert "generateReceptionEventsBulk" in sim and sim.count("{") == sim.count("}")
    # the patched reference classes still call only what exists (incl. what the patch and the adapter add)
    a = api()
    a["CudaRadioMedium"] = dict(methods={n: [["?", -1]] for n in declared}, fields={}, constructors=[], extends="AbstractRadioMedium")
    for rel in ("Simulator.java",):
        p = work / "radio-medium" / "java" / "se" / "sics" / "emul8" / "radiomedium" / rel
        text = strip_comments(open(p).read())
        for m in re.finditer(r"new (TransmissionEvent|BulkReceptionEvent)\(", text):
            close = match_fwd(text, m.end() - 1)
            n = split_args(text[m.end():close])
            assert n == (4 if m.group(1) == "TransmissionEvent" else 10), (m.group(1), n)
