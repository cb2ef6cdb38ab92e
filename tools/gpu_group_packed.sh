export RAFTCORR_LIB=raft_stir_b200/_build/attrib/libraftcorr.so
for g in 3 8 5 3 8 5; do echo "group $g"; RC_TC_GROUP=$g timeout 200 python tools/build_bench.py | grep '"bf16", "math": "bf16"'; done
