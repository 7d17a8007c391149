// tracer_facade.h — the few settings the reference keeps in file-static variables without a getter,
// so the GPU-backed tracer (tracer_facade.cpp) cannot read them back.  A maintainer adds the two
// calls marked in INTEGRATION.md next to the existing ones in src/main.cpp:91-92.
#ifndef PTB_TRACER_FACADE_H
#define PTB_TRACER_FACADE_H
void ptbFacadeSetGridSettings(unsigned size, unsigned oSize, unsigned count); // mirrors setGridSettings, lightFieldGrid.cpp:54-58
void ptbFacadeSetGridBeamsSettings(unsigned size, unsigned count);            // mirrors setGridBeamsSettings, lightFieldGridSubBeams.cpp:179-183
void ptbFacadeSetMaxDepth(int depth);  // 0 = unbounded (reference behaviour); BASELINE.json's "N bounces" = depth N+1
// Mode BVH walks the reference's OWN tree (read through getNode / getNodeAABB / bvhGetTrisIndex) node for node, so that
// getIntersectionCount / getStructureIntersectionCount and the heat maps are the CPU tracer's numbers (slower; off by default)
void ptbFacadeUseReferenceTree(bool on);
void ptbFacadeSetDevice(int device);   // before the first trace(); default 0
// several GPUs of the box behind the one trace() thread: scene replicated, frame tile-sharded (ptb_group_*), same pixels
void ptbFacadeSetDevices(const int *devices, int n);
unsigned long ptbFacadeStructureBytes(int mode); // device bytes of the structure (cf. getMemoryBVH / getMemory2Plane)
double ptbFacadeLastRenderMs();        // device time of the last batch (CUDA events)
unsigned long ptbFacadeRaysTraced();   // traversals executed on the GPU since initTracer/reset
#endif
