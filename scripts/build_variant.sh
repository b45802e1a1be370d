# usage: build_variant.sh <name> "<extra nvcc flags>"   -> scratch/lib_<name>.so
set -e
cd /root/repo
PB_NVCC_EXTRA="$2" python -c "
from pupiloptixlab_b200 import build
build.build_native(force=True)"
cp pupiloptixlab_b200/csrc/libpupil_b200.so scratch/lib_$1.so
grep -A2 "Function properties for _ZN2pb" pupiloptixlab_b200/csrc/pb_wavefront.ptxas.log | grep -E "Function|Used|spill" | paste - - - | grep -E "k_shade|k_extendILb0|k_connectILb0" | sed -E 's/.*(k_[a-z]+IL[ib][0-9]).*frame, ([0-9]+) bytes spill stores, ([0-9]+) bytes spill loads.*Used ([0-9]+) registers.*/\1 regs \4 spill \2\/\3/' | tr '\n' ';'; echo " <- $1"
