#!/bin/bash
# SURVEY Appendix B / VERDICT r1 item 1(a): is anything of the reference's physics stack (MuJoCo, dm_control,
# mujoco_utils, dm_env_wrappers, note_seq, the menagerie Shadow Hand XML) present on the GPU box?
# Raw output goes to gpurun_out/probe_reference_stack.txt; the copy under profiles/ is the committed record.
out=${1:-gpurun_out/probe_reference_stack.txt}
mkdir -p "$(dirname "$out")"
{
echo "== date / host"; date -u; uname -a; nproc; grep -m1 'model name' /proc/cpuinfo
echo "== nvidia-smi"; nvidia-smi --query-gpu=name,driver_version,memory.total --format=csv
echo "== python imports"
for m in mujoco dm_control mujoco_utils dm_env_wrappers dm_env note_seq pretty_midi mido gym gymnasium jax mujoco_mjx mujoco_warp brax sklearn wandb; do
  python - <<PY 2>&1 | tail -1
try:
    import importlib
    m = importlib.import_module("$m")
    print("$m: OK", getattr(m, "__version__", "?"), getattr(m, "__file__", "?"))
except Exception as e:
    print("$m:", type(e).__name__, e)
PY
done
echo "== pip download mujoco (expects: no network)"
timeout 30 python -m pip download --no-deps -d /tmp/_pipdl mujoco 2>&1 | tail -3
echo "== pip index from the offline wheelhouse"
ls /opt/wheelhouse 2>/dev/null | grep -i -E 'mujoco|dm_control|dm-control|dm_env|note_seq|menagerie|labmaze' || echo "(no mujoco / dm_control / note_seq wheel in /opt/wheelhouse)"
echo "== filesystem search"
timeout 120 find / -xdev \( -path /proc -o -path /sys -o -path "$PWD" -o -path /root/repo \) -prune -o \
  \( -iname 'right_hand.xml' -o -iname 'left_hand.xml' -o -iname '*menagerie*' -o -iname 'libmujoco*' -o -iname 'mujoco*' -o -iname 'dm_control*' -o -iname 'shadow_hand*' -o -iname '*distal_pst*' \) -print 2>/dev/null | head -50
echo "(end of search)"
echo "== baseline/_ref"; ls -la baseline/_ref 2>&1 | head
} > "$out" 2>&1
cat "$out"
