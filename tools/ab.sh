source tools/variants.sh true
run w18
run w20 -DEPLAN_WARPS=20
run w17 -DEPLAN_WARPS=17
