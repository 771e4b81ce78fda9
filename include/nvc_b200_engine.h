/* placeholder: filled in below */
#ifndef NVC_B200_ENGINE_H
#define NVC_B200_ENGINE_H
#include "nvc_b200.h"
#endif
