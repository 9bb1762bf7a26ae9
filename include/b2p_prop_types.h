/* Collider classes of the reference's props beyond the box (mujoco_robot_environments/environment/props.py): exactly what each
 * class's MJCF sets on its collision geom; every other field is MuJoCo's default.  Shared by the kernels (csrc/) and the oracle.
 *
 *   Cylinder   props.py:307-388   type="cylinder" size=(radius, half_height) mass=0.1 condim=6 solref=(0.01, 1.1) friction=(1,1,1) priority=10
 *   GalaApple  props.py:463-489 + assets/mjcf/gala_apple/Apple.xml   <default class="collision"> type="sphere" size="0.02": all defaults
 *              (condim 3, friction 1 / 0.005 / 0.0001, priority 0, solref 0.02 / 1).  The body's mass comes from its geoms at density
 *              1000; the visual mesh Apple_edit.obj the XML names is NOT in the reference tree, so only the sphere contributes here.
 *   Rectangle  props.py:171-304   the box class: the compiled blob's prop geoms carry it (priority 10, margin = gap = 0.15), no entry here.
 *
 * The prop's class is per-scene data (add_objects draws a shape per prop, props.py:604-617): b2p_set_env_prop_types. */
#ifndef B2P_PROP_TYPES_H
#define B2P_PROP_TYPES_H
#include "b2p_model_layout.h"

typedef struct {
  int geom_type, condim, priority;
  double friction[3], solref[2], solimp[5], solmix, margin, gap;
  double mass;      /* > 0: as given; < 0: density 1000 x volume */
} b2p_prop_class;

static const b2p_prop_class B2P_PROP_CLASS_SPHERE = {B2P_GEOM_SPHERE, 3, 0, {1.0, 0.005, 0.0001}, {0.02, 1.0}, {0.9, 0.95, 0.001, 0.5, 2.0}, 1.0, 0.0, 0.0, -1.0};
static const b2p_prop_class B2P_PROP_CLASS_CYLINDER = {B2P_GEOM_CYLINDER, 6, 10, {1.0, 1.0, 1.0}, {0.01, 1.1}, {0.9, 0.95, 0.001, 0.5, 2.0}, 1.0, 0.0, 0.0, 0.1};

/* mass, principal inertia and bounding half extents of a prop of the given class and size (sphere: size[0] = radius; cylinder:
 * size[0] = radius, size[1] = half height, axis = local z; box: half extents).  box_mass: the blob's prop mass. */
static inline void b2p_prop_inertia(int geom_type, const double* size, double box_mass, double* mass, double* inertia, double* half) {
  const double pi = 3.14159265358979323846;
  if (geom_type == B2P_GEOM_SPHERE) {
    const double r = size[0], m = B2P_PROP_CLASS_SPHERE.mass > 0 ? B2P_PROP_CLASS_SPHERE.mass : 1000.0 * 4.0 / 3.0 * pi * r * r * r;
    *mass = m; inertia[0] = inertia[1] = inertia[2] = 0.4 * m * r * r; half[0] = half[1] = half[2] = r;
  } else if (geom_type == B2P_GEOM_CYLINDER) {
    const double r = size[0], h = size[1], m = B2P_PROP_CLASS_CYLINDER.mass > 0 ? B2P_PROP_CLASS_CYLINDER.mass : 1000.0 * pi * r * r * 2 * h;
    *mass = m; inertia[0] = inertia[1] = m * (0.25 * r * r + h * h / 3.0); inertia[2] = 0.5 * m * r * r; half[0] = half[1] = r; half[2] = h;
  } else {
    const double m = box_mass;
    *mass = m; inertia[0] = m / 3 * (size[1] * size[1] + size[2] * size[2]); inertia[1] = m / 3 * (size[0] * size[0] + size[2] * size[2]);
    inertia[2] = m / 3 * (size[0] * size[0] + size[1] * size[1]); half[0] = size[0]; half[1] = size[1]; half[2] = size[2];
  }
}
#endif
