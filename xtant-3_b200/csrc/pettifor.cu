// Pettifor (orthogonal sp3) branch -- placeholder until the kernels land.
#include "xtb_ctx.h"
int xtb_impl_pettifor_assemble(xtb_ctx* c) { XTB_FAIL(c, XTB_EINVAL, "Pettifor assembly is not built yet"); }
int xtb_impl_pettifor_forces(xtb_ctx* c) { XTB_FAIL(c, XTB_EINVAL, "Pettifor forces are not built yet"); }
int xtb_impl_pettifor_pressure(xtb_ctx* c) { XTB_FAIL(c, XTB_EINVAL, "Pettifor pressure is not built yet"); }
