// Boundary classification of a mesh node; same enumerators and values as the reference (src/NodeType.h:4-8),
// because the values travel through mask files (Experiments/Results/BaseCircle/*mask.txt) and the C ABI.
#ifndef NODE_TYPE_H
#define NODE_TYPE_H
enum NodeType { BOUNDARY_FREE = 0, BOUNDARY_FIXED = 1, INTERIOR = 2 };
#endif
