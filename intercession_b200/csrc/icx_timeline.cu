// World-line record / replay / divergence scan.  (filled in below)
#include "icx_host.h"
