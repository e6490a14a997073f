#!/bin/bash
# usage: scripts/role_prof.sh <prof script>  -- runs a -DVK_ROLE_PROF build of the library; prints per-role cycle counters
python "$1" 2>&1 | sort | uniq -c | sort -rn | head -40
