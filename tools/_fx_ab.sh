for v in "" co_s2 w8s2 ""; do
  if [ -z "$v" ]; then unset RTB_LIB_PATH; else export RTB_LIB_PATH=build/variants/lib_$v.so; fi
  echo "== ${v:-new}" ; python tools/ab_probe.py 0 0 60 2>&1 | tail -2 | cut -c1-330
done
