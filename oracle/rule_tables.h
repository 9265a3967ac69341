/* TEST INFRASTRUCTURE: the quadrature tables are data dumped from the compiled reference (tools/gen_rule_tables.py);
 * one copy lives beside the kernels, the oracle reads the same file. */
#include "../ngsxfem_b200/csrc/rule_tables.h"
