#!/usr/bin/env python
"""Extract one kernel's SASS from `cuobjdump -sass <obj>` (encodings stripped): python tools/sass_extract.py obj mangled_name"""
import re
import subprocess
import sys

txt = subprocess.run(["cuobjdump", "-sass", sys.argv[1]], capture_output=True, text=True).stdout.splitlines()
on = False
for line in txt:
    if "Function :" in line:
        on = sys.argv[2] in line
    if on and not re.match(r"^\s*/\* 0x[0-9a-f]+ \*/\s*$", line):
        print(re.sub(r"\s*/\* 0x[0-9a-f]+ \*/\s*$", "", line))
