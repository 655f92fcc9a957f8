// writewatch.h -- write tracking of caller-owned host arrays (writewatch.cpp): userfaultfd async
// write-protection + PAGEMAP_SCAN.  Host only.  Everything answers "cannot tell" where the kernel
// lacks the facility; callers then hash.
#pragma once

#include <stddef.h>

namespace cgb {

bool ww_available();                          // facility present and self-tested (CGB200_WRITE_TRACKING=0: off)
bool ww_arm(const void* p, size_t bytes);     // start watching the page-aligned interior of [p, p+bytes)
bool ww_rearm(const void* p, size_t bytes);   // forget the writes seen so far
int ww_written(const void* p, size_t bytes);  // 1 = written since armed, 0 = not, -1 = cannot tell
void ww_disarm(const void* p, size_t bytes);
size_t ww_page();

}  // namespace cgb
