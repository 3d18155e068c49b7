"""Print the SASS of one kernel of libh3d.so (substring match on the mangled name)."""
import re, subprocess, sys
so = "pointcloud_segmentation_b200/lib/libh3d.so"
pat = sys.argv[1]
txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
blocks = re.split(r"\n\s*Function : ", txt)
for b in blocks[1:]:
    name = b.split("\n", 1)[0].strip()
    if pat in name:
        print("Function :", name)
        for line in b.split("\n"):
            m = re.match(r"\s+/\*([0-9a-f]{4,6})\*/\s+(.*?);", line)
            if m:
                print(m.group(1), m.group(2))
        break
