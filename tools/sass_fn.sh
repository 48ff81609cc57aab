#!/bin/bash
# usage: tools/sass_fn.sh <object> <substring of the mangled name>  -- prints the SASS instructions of the first match
cuobjdump -sass "$1" 2>/dev/null | awk -v k="$2" '/Function : /{ if (f) exit; f = index($0, k) > 0 } f' \
  | grep -E "^\s+/\*[0-9a-f]{4,5}\*/" | sed 's#/\* 0x[0-9a-f]* \*/##'
