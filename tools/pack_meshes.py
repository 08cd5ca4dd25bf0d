"""Packs the reference's mesh input files into the binary .s35m container under meshes/.

The GPU box has no /root/reference, so the three meshes every benchmark configuration starts from travel
with the repository as data: vertices, cell vertex ids (lexicographic) and boundary faces with their
indicators, exactly what the .inp / .msh readers of the host library produce.  Run from the repo root:
    python tools/pack_meshes.py [/root/reference]
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402


def main():
    ref = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
    pkg = ge.load_package()
    out_dir = os.path.join(ROOT, "meshes")
    os.makedirs(out_dir, exist_ok=True)
    for name, dim in (("nsbench2.inp", 2), ("cylinder3d.msh", 3), ("cylinderextruded.msh", 3)):
        out = os.path.join(out_dir, os.path.splitext(name)[0] + ".s35m")
        pkg.mesh_convert(os.path.join(ref, name), dim, out)
        print("%s -> %s (%d bytes)" % (name, out, os.path.getsize(out)))


if __name__ == "__main__":
    main()
