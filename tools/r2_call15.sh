#!/bin/bash
bash tools/ab_env.sh "" "AY_STAGE_KB=64" "" "AY_STAGE_KB=64"
O=gpurun_out/ab
echo "layer | default | STAGE_KB=64"
paste <(grep -E "^ *(63|66|69|87|99|38|13) " $O/lt_env1.txt | cut -c1-44) <(grep -E "^ *(63|66|69|87|99|38|13) " $O/lt_env2.txt | cut -c33-44)
