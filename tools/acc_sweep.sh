#!/bin/bash
# calibration of the truncation compensation for one engine: rel_err per shape under several OCPPO_TC_TRUNC values
for t in ${TRUNCS:-0 0.5 1 1.5 2}; do echo "== trunc_scale $t"; OCPPO_TC_TRUNC=$t bash tools/acc.sh ${ENG:-8} 2>&1 | cut -c1-100; done
