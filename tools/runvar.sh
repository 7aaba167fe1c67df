python tools/time_variants.py 2>&1 | tail -1
for f in build/var/lib_*.so; do echo $f; MG_LIB_PATH=$PWD/$f python tools/time_variants.py 2>&1 | tail -1; done
python tools/time_variants.py FindDupe-TestLayout-LoRes4E-v0 16384 2>&1 | tail -1
python tools/time_variants.py ClusterColour-TestAll-LoRes4E-v0 16384 2>&1 | tail -1
python tools/time_variants.py MoveToCorner-Demo-LoRes4E-v0 16384 2>&1 | tail -1
