// facade_internal.h — what the two facade translation units share (not part of the reference's interface).
// tracer_facade.cpp owns the device contexts; structures_facade.cpp fronts the structure headers with them.
#ifndef PTB_FACADE_INTERNAL_H
#define PTB_FACADE_INTERNAL_H
#include "ptb.h"
namespace ptbfacade {
ptb_ctx *context();                 // member 0 of the device group, created on first use; nullptr if that failed
bool sceneOnDevice();               // getTris()/getObjects()/getMaterials() uploaded and the BVH built (again if the triangle list changed)
bool structureOnDevice(int mode);   // + the structure mode needs (TWO_PLANE: Subgrids, SUB_BEAMS: beams), with the current settings
void structuresAreFronted();        // structures_facade.cpp is linked: the reference's own tree does not exist in this process
}
#endif
