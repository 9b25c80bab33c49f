#!/bin/bash
# A/B of library variants over the headline scene, the 1 M-triangle mesh and the generic-tier scenes:
#   bash tests/tools/ab_all.sh "" _l0      (variant suffixes of raysnail_b200/libraysnail_b200<suffix>.so)
for v in "$@"; do
  bash profiles/quick.sh "$v"
  bash tests/tools/mesh_ab.sh "variant[$v]||$v"
  SCENES="${SCENES:-quadric csg cornell example}" bash tests/tools/ab_scenes.sh "variant[$v]||$v"
done
