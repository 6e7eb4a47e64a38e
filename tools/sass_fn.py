#!/usr/bin/env python
"""Print the SASS of the functions whose mangled name contains every given substring.
usage: cuobjdump -sass lib.so | python tools/sass_fn.py ILb0ELi3 LinearFieldELb0ELi256"""
import sys
keys = sys.argv[1:]
on = False
for line in sys.stdin:
    if "Function :" in line:
        on = all(k in line for k in keys)
    if on:
        sys.stdout.write(line)
