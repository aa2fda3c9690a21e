# value-only bench for every library under .variants/ (and the default build)
for L in "" $(ls .variants/*.so 2>/dev/null); do echo "== ${L:-default}"; ORBX_LIB=${L:+$PWD/$L} timeout 120 python bench.py --steps 20 --warmup 3 --value-only 2>/dev/null | tail -1 | python -c "import json,sys; j=json.loads(sys.stdin.read()); print(round(j['value']), round(j['ms_per_step'],3))"; done
