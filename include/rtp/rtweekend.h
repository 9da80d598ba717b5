// rtweekend.h -- same name as the reference header (src/rtweekend.h); the API lives in scene_api.h
#include "scene_api.h"
