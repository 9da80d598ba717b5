// bvh.h -- same name as the reference header (src/bvh.h); the API lives in scene_api.h
#include "scene_api.h"
