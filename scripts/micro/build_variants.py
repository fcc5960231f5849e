"""Build kernel variants of libtoric_b200.so for on-GPU A/B timing (scripts/micro/variants/*.so)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from toric_rl_decoder_b200 import buildlib
VAR = {
    "g0_w4_s1_t2": ["TC_WRITE_WAVES=0", "TC_WRITE_WARPS=4", "TC_STAGE_BUFS=1", "TC_TILE_WPL=2"],
    "g0_w4_s2_t1": ["TC_WRITE_WAVES=0", "TC_WRITE_WARPS=4", "TC_STAGE_BUFS=2", "TC_TILE_WPL=1"],
    "g0_w2_s2_t1": ["TC_WRITE_WAVES=0", "TC_WRITE_WARPS=2", "TC_STAGE_BUFS=2", "TC_TILE_WPL=1"],
    "g0_w2_s1_t2": ["TC_WRITE_WAVES=0", "TC_WRITE_WARPS=2", "TC_STAGE_BUFS=1", "TC_TILE_WPL=2"],
    "g0_w4_s3_t1": ["TC_WRITE_WAVES=0", "TC_WRITE_WARPS=4", "TC_STAGE_BUFS=3", "TC_TILE_WPL=1"],
    "g0_w1_s2_t2": ["TC_WRITE_WAVES=0", "TC_WRITE_WARPS=1", "TC_STAGE_BUFS=2", "TC_TILE_WPL=2"],
    "g4_w4_s1_t2": ["TC_WRITE_WAVES=4", "TC_WRITE_WARPS=4", "TC_STAGE_BUFS=1", "TC_TILE_WPL=2"],
    "g4_w4_s3_t1": ["TC_WRITE_WAVES=4", "TC_WRITE_WARPS=4", "TC_STAGE_BUFS=3", "TC_TILE_WPL=1"],
    "g0_w4_s2_t2": ["TC_WRITE_WAVES=0", "TC_WRITE_WARPS=4", "TC_STAGE_BUFS=2", "TC_TILE_WPL=2"],
}
out_dir = os.path.join(ROOT, "scripts", "micro", "variants")
os.makedirs(out_dir, exist_ok=True)
for name, defs in VAR.items():
    if len(sys.argv) > 1 and name not in sys.argv[1:]:
        continue
    print(buildlib.build(defines=defs, out=os.path.join(out_dir, "lib_%s.so" % name)))
