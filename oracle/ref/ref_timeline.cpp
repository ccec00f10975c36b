// TEST INFRASTRUCTURE -- not product code (see ref_harness.cpp header).
// Reference timestream containers (EntityTimestreamMap / TsBreakpointQueue /
// Message / coherency compare), header-only in the reference, driven headlessly.
#include "../../include/icx.h"
extern "C" int icxref_timeline_placeholder(void) { return 0; }
