// aabb.h -- same name as the reference header (src/aabb.h); the API lives in scene_api.h
#include "scene_api.h"
