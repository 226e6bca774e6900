"""Golden vector of the rig_calibrator command line, taken from the reference's own source in THIS container
(/root/reference does not travel to the GPU box): the name, type and default of every DEFINE_* flag of
rig_calibrator/bin/rig_calibrator.cc in declaration order, and the text of every LOG(FATAL) of parameterValidation().
Writes tests/golden/cli_flags.json.  Run: python tools/make_golden_cli.py"""
import json
import os
import re

SRC = "/root/reference/rig_calibrator/bin/rig_calibrator.cc"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
text = open(SRC).read()

flags = []
for m in re.finditer(r"DEFINE_(double|int32|string|bool)\(\s*(\w+)\s*,\s*([^,]+?)\s*,", text):
    kind, name, default = m.groups()
    default = default.strip()
    if kind == "string":
        default = json.loads(default)
    elif kind == "bool":
        default = default == "true"
    elif kind == "int32":
        default = int(default)
    else:
        default = float(default)
    flags.append({"name": name, "type": {"double": "d", "int32": "i", "string": "s", "bool": "b"}[kind], "default": default})

body = text[text.index("void parameterValidation()"):text.index("void set_up_block_sizes")]
messages = []
for m in re.finditer(r"LOG\(FATAL\)\s*<<(.*?);", body, re.S):
    parts = re.findall(r'"((?:[^"\\]|\\.)*)"', m.group(1))
    messages.append("".join(json.loads('"%s"' % p) for p in parts))

out = {"source": "rig_calibrator/bin/rig_calibrator.cc DEFINE_* (:135-316) and parameterValidation (:893-955)", "flags": flags,
       "validation_messages": messages}
with open(os.path.join(ROOT, "tests", "golden", "cli_flags.json"), "w") as f:
    json.dump(out, f, indent=1)
print(len(flags), "flags,", len(messages), "messages")
