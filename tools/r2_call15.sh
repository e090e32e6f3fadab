#!/bin/bash
bash tools/ab_env.sh "" "AY_RES_KB=128" "" "AY_RES_KB=128"
O=gpurun_out/ab
echo "layer | default | RES_KB=128"
paste <(grep -E "^ *(38|41|44|47|50|53|56|59|89|91|87|96) " $O/lt_env1.txt | cut -c1-44) <(grep -E "^ *(38|41|44|47|50|53|56|59|89|91|87|96) " $O/lt_env2.txt | cut -c33-44)
