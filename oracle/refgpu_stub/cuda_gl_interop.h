/* oracle/refgpu_stub/cuda_gl_interop.h -- TEST INFRASTRUCTURE ONLY.
 * Empty shadow of the CUDA toolkit header of the same name, placed FIRST on the include path when
 * oracle/Makefile compiles the reference's GpuPathTracer/tracer.cu in place (tracer.cu:17 includes it; the real
 * header needs GL/gl.h, and there is no graphics stack on this box).  The kernel uses nothing from it. */
