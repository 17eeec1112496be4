"""Build step of oracle/_ref/*.so (oracle/Makefile): writes, into a build intermediate that the Makefile deletes after the
compile, the DEFINITIONS of a few functions and tables of the reference, each cut from the file where it lies — from the first
line of the definition to its closing line — so that they can be compiled against the stub class declarations of
oracle/ref_stubs/ and oracle/ref_app/.  (The files as a whole — src/base/dof_map_constraints.C, 5 000 lines over the whole
DofMap; src/fe/fe_boundary.C and h1_fe_transformation.C with their instantiations for every family; the element classes of
src/geom/ — cannot be compiled stand-alone.)  Nothing is committed: the output goes under oracle/_ref/, which is git-ignored.

    python3 ref_excerpt.py <reference root> <set: fe | app | mesh_idx | mesh_so | mesh> <output file>
"""
import re
import sys

ref, which, out = sys.argv[1], sys.argv[2], sys.argv[3]
FUNC_END, TABLE_END = r"^}$", r"^  };$"
SETS = {
    # (file, first line of the definition, last line, text put in front)
    "fe": [
        ("src/fe/fe_boundary.C", r"^void FEMap::init_face_shape_functions\(", FUNC_END, "template<unsigned int Dim>\n"),
        ("src/fe/fe_boundary.C", r"^void FEMap::compute_face_map\(", FUNC_END, ""),
        ("src/fe/fe_abstract.C", r"^void FEAbstract::get_refspace_nodes\(", FUNC_END, ""),
        ("src/geom/cell_hex8.C", r"^const unsigned int Hex8::side_nodes_map\[", TABLE_END, ""),
        ("src/geom/cell_hex20.C", r"^const unsigned int Hex20::side_nodes_map\[", TABLE_END, ""),
        ("src/geom/cell_hex27.C", r"^const unsigned int Hex27::side_nodes_map\[", TABLE_END, ""),
        ("src/geom/cell_tet10.C", r"^const unsigned int Tet10::side_nodes_map\[", TABLE_END, ""),
        ("src/geom/cell_prism18.C", r"^const unsigned int Prism18::side_nodes_map\[", TABLE_END, ""),
        ("src/geom/cell_hex8.C", r"^bool Hex8::has_affine_map\(\) const", FUNC_END, ""),
        ("src/geom/cell_hex20.C", r"^bool Hex20::has_affine_map\(\) const", FUNC_END, ""),
        ("src/geom/cell_hex27.C", r"^bool Hex27::has_affine_map\(\) const", FUNC_END, ""),
        ("src/geom/cell_tet10.C", r"^bool Tet10::has_affine_map\(\) const", FUNC_END, ""),
        ("src/geom/cell_prism18.C", r"^bool Prism18::has_affine_map\(\) const", FUNC_END, ""),
    ],
    "mesh_idx": [
        ("src/mesh/mesh_generation.C", r"^unsigned int idx\(const ElemType type,", FUNC_END, "inline\n"),
    ],
    "mesh_so": [
        ("src/mesh/unstructured_mesh.C", r"^map_hi_order_node\(unsigned int hon,", FUNC_END, "std::map<std::vector<dof_id_type>, Node *>::iterator\n"),
        ("src/mesh/unstructured_mesh.C", r"^void transfer_elem\(Elem & lo_elem,", FUNC_END, ""),
        ("src/mesh/unstructured_mesh.C", r"^all_increased_order_range \(UnstructuredMesh & mesh,", FUNC_END, "template <typename ElemTypeConverter>\nvoid\n"),
    ],
    "mesh": [
        ("src/mesh/mesh_generation.C", r"^void MeshTools::Generation::build_cube\(", FUNC_END, ""),
        ("src/geom/cell_hex27.C", r"^const unsigned int Hex27::side_nodes_map\[", TABLE_END, ""),
        ("src/mesh/replicated_mesh.C", r"^void ReplicatedMesh::renumber_nodes_and_elements \(\)", FUNC_END, ""),
        ("src/base/dof_map.C", r"^void DofMap::distribute_local_dofs_var_major$", FUNC_END, ""),
        ("src/partitioning/linear_partitioner.C", r"^void LinearPartitioner::partition_range\(", FUNC_END, ""),
        ("src/mesh/unstructured_mesh.C", r"^UnstructuredMesh::all_second_order_range \(", FUNC_END, "void\n"),
        ("src/geom/elem.C", r"^ElemType Elem::second_order_equivalent_type \(const ElemType et,", FUNC_END, ""),
        ("src/geom/cell_tet10.C", r"^const unsigned short int Tet10::_second_order_adjacent_vertices\[6\]\[2\] =", TABLE_END, ""),
    ],
    "app": [
        ("src/base/dof_map_constraints.C", r"^void DofMap::build_constraint_matrix \(", FUNC_END, ""),
        ("src/base/dof_map_constraints.C", r"^void DofMap::build_constraint_matrix_and_vector \(", FUNC_END, ""),
        ("src/base/dof_map_constraints.C", r"^void DofMap::constrain_element_matrix_and_vector \(", FUNC_END, ""),
        ("src/base/dof_map_constraints.C", r"^void DofMap::heterogeneously_constrain_element_matrix_and_vector \(", FUNC_END, ""),
        ("src/fe/h1_fe_transformation.C", r"^void H1FETransformation<OutputShape>::map_dphi\(", FUNC_END, "template<typename OutputShape>\n"),
    ],
}
with open(out, "w") as o:
    for path, start, stop, prefix in SETS[which]:
        lines = open(f"{ref}/{path}").read().split("\n")
        hits = [i for i, l in enumerate(lines) if re.match(start, l)]
        if len(hits) != (2 if which == "mesh_idx" else 1):  # idx() has a 2D and a 3D overload
            sys.exit(f"ref_excerpt: {len(hits)} definitions match {start!r} in {path}")
        for h in hits:
            end = next(i for i in range(h, len(lines)) if re.match(stop, lines[i]))
            o.write(f"// ---- {path}:{h + 1}-{end + 1}\n{prefix}")
            o.write("\n".join(lines[h:end + 1]) + "\n\n")
