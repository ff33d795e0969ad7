source tools/variants.sh true
run new
run new16 -DEPLAN_WARPS=16
runref old
