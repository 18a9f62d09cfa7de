/* Forwarder: the declarations a caller expects from "amcl/pf/pf_pdf.h" live in one
 * combined ABI header (see its preamble for the reference citations). */
#include "../../amcl_b200/amcl_abi.h"
