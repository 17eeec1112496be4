"""Build step of oracle/_ref/libref_app.so (oracle/Makefile): writes, into a build intermediate that the Makefile deletes after
the compile, the DEFINITIONS of a few member functions of the reference, each cut from the file where it lies — from the first
line of the definition to its closing brace at column 0 — so that they can be compiled against the stub class declarations of
oracle/ref_app/ref_world.h.  (The files as a whole — src/base/dof_map_constraints.C, 5 000 lines over the whole DofMap, and
src/fe/h1_fe_transformation.C with its vector-valued instantiations — cannot be compiled stand-alone.)  Nothing is committed:
the output goes under oracle/_ref/, which is git-ignored."""
import re
import sys

ref, out = sys.argv[1], sys.argv[2]
WANT = [
    ("src/base/dof_map_constraints.C", r"^void DofMap::build_constraint_matrix \(", ""),
    ("src/base/dof_map_constraints.C", r"^void DofMap::build_constraint_matrix_and_vector \(", ""),
    ("src/base/dof_map_constraints.C", r"^void DofMap::constrain_element_matrix_and_vector \(", ""),
    ("src/base/dof_map_constraints.C", r"^void DofMap::heterogeneously_constrain_element_matrix_and_vector \(", ""),
    ("src/fe/h1_fe_transformation.C", r"^void H1FETransformation<OutputShape>::map_dphi\(", "template<typename OutputShape>\n"),
]
with open(out, "w") as o:
    for path, start, prefix in WANT:
        lines = open(f"{ref}/{path}").read().split("\n")
        hits = [i for i, l in enumerate(lines) if re.match(start, l)]
        if len(hits) != 1:
            sys.exit(f"ref_excerpt: {len(hits)} definitions match {start!r} in {path}")
        end = next(i for i in range(hits[0], len(lines)) if lines[i] == "}")
        o.write(f"// ---- {path}:{hits[0] + 1}-{end + 1}\n{prefix}")
        o.write("\n".join(lines[hits[0]:end + 1]) + "\n\n")
