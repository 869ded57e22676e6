"""Generates tests/golden/reference_java_api.json: the public methods (name -> [return type, number of parameters])
and public fields of the reference classes the Java adapter touches, parsed from the reference SOURCE
(/root/reference/radio-medium/java).  No JDK exists in this image, so this is how tests/test_java_binding.py knows
that `packet.getTransmitPower()` exists and `packet.getTxpower()` does not.

    python tests/golden/make_reference_java_api.py
"""
import json
import os
import re

REF = "/root/reference/radio-medium/java"
CLASSES = {
    "Simulator": "se/sics/emul8/radiomedium/Simulator.java",
    "Node": "se/sics/emul8/radiomedium/Node.java",
    "Transciever": "se/sics/emul8/radiomedium/Transciever.java",
    "RadioPacket": "se/sics/emul8/radiomedium/RadioPacket.java",
    "Position": "se/sics/emul8/radiomedium/Position.java",
    "RadioMedium": "se/sics/emul8/radiomedium/RadioMedium.java",
    "AbstractRadioMedium": "se/sics/emul8/radiomedium/AbstractRadioMedium.java",
    "UDGMRadioMedium": "se/sics/emul8/radiomedium/UDGMRadioMedium.java",
    "N2NRadioMedium": "se/sics/emul8/radiomedium/N2NRadioMedium.java",
    "NullRadioMedium": "se/sics/emul8/radiomedium/NullRadioMedium.java",
    "UDGMConstantLossRadioMedium": "se/sics/emul8/radiomedium/UDGMConstantLossRadioMedium.java",
    "TimeEvent": "com/botbox/scheduler/TimeEvent.java",
    "EventQueue": "com/botbox/scheduler/EventQueue.java",
    "TransmissionEvent": "se/sics/emul8/radiomedium/events/TransmissionEvent.java",
    "ReceptionEvent": "se/sics/emul8/radiomedium/events/ReceptionEvent.java",
}
METHOD = re.compile(r"^\s*(?:public|protected)\s+(?:static\s+|final\s+|synchronized\s+|abstract\s+)*([\w<>\[\],.]+)\s+(\w+)\s*\(([^)]*)\)", re.M)
CTOR = re.compile(r"^\s*public\s+(\w+)\s*\(([^)]*)\)", re.M)
FIELD = re.compile(r"^\s*(?:public|protected)\s+(?:static\s+|final\s+)*([\w<>\[\]]+)\s+(\w+)\s*(?:=[^;]*)?;", re.M)


def strip_comments(text):
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return re.sub(r"//[^\n]*", "", text)


def nargs(params):
    params = params.strip()
    return 0 if not params else params.count(",") + 1


def parse(path):
    text = strip_comments(open(path).read())
    methods, fields, ctors = {}, {}, []
    for ret, name, params in METHOD.findall(text):
        methods.setdefault(name, []).append([ret, nargs(params)])
    for name, params in CTOR.findall(text):
        ctors.append(nargs(params))
    for typ, name in FIELD.findall(text):
        fields[name] = typ
    ext = re.search(r"class\s+\w+\s+extends\s+(\w+)", text)
    return dict(methods=methods, fields=fields, constructors=sorted(set(ctors)), extends=ext.group(1) if ext else None)


def build():
    return {cls: parse(os.path.join(REF, rel)) for cls, rel in CLASSES.items()}


if __name__ == "__main__":
    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "reference_java_api.json")
    with open(out, "w") as f:
        json.dump(build(), f, indent=1, sort_keys=True)
        f.write("\n")
    print("wrote", out)
