#!/bin/bash
# SASS opcode evidence for the built objects (no GPU needed): tcgen05 MMA (UTCHMMA), TMA loads (UTMALDG) and
# descriptor prefetch (UTMACCTL.PF), TMEM loads (LDTM), tcgen05 commit barriers (UTCBAR), TMEM alloc (UTCATOMSWS /
# UTCHMMA companions), programmatic dependent launch (PREEXIT = griddepcontrol.launch_dependents,
# ACQBULK = griddepcontrol.wait).  Usage: bash tools/sass_table.sh > profiles/r2_sass_opcodes.txt
echo "# cuobjdump -sass of minpy_b200/_build/*.o (sm_100a), opcode counts per object; $(date -u +%Y-%m-%dT%H:%MZ); nvcc $(nvcc --version | tail -1)"
printf "%-16s %8s %8s %8s %8s %8s %8s %8s %8s %8s\n" object UTCHMMA UTMALDG UTMACCTL LDTM UTCBAR UTMASTG PREEXIT ACQBULK kernels
for o in minpy_b200/_build/*.o; do
  s=$(cuobjdump -sass $o 2>/dev/null)
  c() { echo "$s" | grep -c "$1"; }
  printf "%-16s %8d %8d %8d %8d %8d %8d %8d %8d %8d\n" $(basename $o) $(c "UTCHMMA") $(c "UTMALDG") $(c "UTMACCTL") $(c " LDTM") $(c "UTCBAR") $(c "UTMASTG") $(c "PREEXIT") $(c "ACQBULK") $(c "Function :")
done
echo
echo "# per tcgen05 kernel"
for o in minpy_b200/_build/gemm.o minpy_b200/_build/conv_tc.o minpy_b200/_build/lstm_persist.o; do
  cuobjdump -sass $o 2>/dev/null | awk '/Function :/{name=$3} /UTCHMMA/{m[name]++} /UTMALDG/{t[name]++} / LDTM/{l[name]++} END{for(k in m) printf "%-110s UTCHMMA %3d UTMALDG %3d LDTM %3d\n", k, m[k], t[k], l[k]}' | sort
done
