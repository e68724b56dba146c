"""Regenerate orb-slam3-noted_b200/csrc/orb_pattern.inc from the reference's table.

TEST INFRASTRUCTURE / provenance script: runs only where /root/reference is mounted.  The table is
the public OpenCV ORB rBRIEF pattern that ORBextractor.cc:148-406 embeds as bit_pattern_31_; it is
data the descriptor must share bit-for-bit, not code.
"""
import re
import sys

REF = "/root/reference/src/ORBextractor.cc"
OUT = "orb-slam3-noted_b200/csrc/orb_pattern.inc"


def main():
    lines = open(REF).read().split("\n")[147:406]
    txt = re.sub(r"/\*.*?\*/", "", "\n".join(lines))
    nums = [int(x) for x in re.findall(r"-?\d+", txt.split("{", 1)[1])]
    assert len(nums) == 1024, len(nums)
    out = [
        "/* rBRIEF learned sampling pattern for 31x31 patches (the public OpenCV ORB table that",
        " * ORBextractor.cc:148-406 carries as bit_pattern_31_): 256 tests x (x0,y0,x1,y1), int8.",
        " * Data only; regenerate with oracle/py/make_pattern.py. */",
    ]
    for i in range(0, 1024, 32):
        out.append(",".join(str(v) for v in nums[i:i + 32]) + ",")
    open(OUT, "w").write("\n".join(out) + "\n")


if __name__ == "__main__":
    sys.exit(main())
