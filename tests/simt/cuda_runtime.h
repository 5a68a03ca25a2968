// Found first on the include path of the host-emulation build: replaces the CUDA runtime header.
#pragma once
#include "simt_host.h"
