source tools/variants.sh true
run c2
run c1 -DEPLAN_CUNROLL=1
run c4 -DEPLAN_CUNROLL=4
runref old
