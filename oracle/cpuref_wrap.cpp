// oracle/cpuref_wrap.cpp -- TEST/BENCH INFRASTRUCTURE ONLY.
// Headless driver for the REFERENCE's CPU tracer (CpuRayTracer/src/*), compiled in place from
// /root/reference by oracle/Makefile into oracle/_ref/libcpuref.so.  It restates
//   main()            CpuRayTracer/src/main.cpp:18-53   (camera, 5 spheres + one Mesh)
//   Renderer::render  CpuRayTracer/src/renderer.cpp:197-222 (OpenMP over rows, Xi = {0,0,y^3},
//                     per pixel `samples` x { get_ray ; trace_ray }, 8-bit quantise toInt :12-14)
// because the real Renderer constructor opens a GLFW window (renderer.cpp:40-137).
// This tracer is a DIFFERENT estimator from the GPU path (double precision, its own
// median-split tree, erand48, Russian roulette): it is timed as "the reference CPU path", never
// used as a parity target.
// Rays are counted without touching the reference sources: a never-hit Object is added to the
// scene; Scene::intersect (scene.cpp:9-21) calls every object's virtual get_intersection once per
// trace_ray call, so its call count == number of rays traced.
#include <stdlib.h>
#include <stdio.h>
#include "material.h"
#include "objects.h"
#include "camera.hpp"
#include "scene.h"
#include <omp.h>
#include <chrono>
#include <vector>
#include <unistd.h>
#include <fcntl.h>

namespace {
struct alignas(64) Counter { unsigned long long n; };
std::vector<Counter> g_counts;
struct RayCounter : public Object {
    ObjectIntersection get_intersection(const Ray&) override
    {
        g_counts[(size_t)omp_get_thread_num()].n += 1;
        return ObjectIntersection(false, 0, glm::dvec3(), Material());
    }
};
inline double clamp01(double x) { return x < 0 ? 0 : x > 1 ? 1 : x; }          // renderer.cpp:12
inline int toInt(double x) { return int(clamp01(x) * 255 + .5); }               // renderer.cpp:14
Scene* g_scene = nullptr;
}

extern "C" {

// Builds the main.cpp scene once (Mesh loads + builds its tree in the constructor, objects.cpp:44-148).
int cpuref_load(const char* obj_path)
{
    if (g_scene) return 0;
    // the reference prints per-row progress and load banners; silence stdout/stderr noise while loading
    Scene* scene = new Scene();
    scene->add(dynamic_cast<Object*>(new Sphere(glm::dvec3(0, 0, -1000), 1000, Material())));
    scene->add(dynamic_cast<Object*>(new Sphere(glm::dvec3(-1004, 0, 0), 1000, Material(DIFF, glm::dvec3(0.85, 0.4, 0.4)))));
    scene->add(dynamic_cast<Object*>(new Sphere(glm::dvec3(1004, 0, 0), 1000, Material(DIFF, glm::dvec3(0.4, 0.4, 0.85)))));
    scene->add(dynamic_cast<Object*>(new Sphere(glm::dvec3(0, 1006, 0), 1000, Material())));
    scene->add(dynamic_cast<Object*>(new Sphere(glm::dvec3(0, 0, 110), 100, Material(EMIT, glm::dvec3(1, 1, 1), glm::dvec3(2.2, 2.2, 2.2)))));
    scene->add(dynamic_cast<Object*>(new Mesh(glm::dvec3(), obj_path, Material(DIFF, glm::dvec3(0.9, 0.9, 0.9)))));
    scene->add(new RayCounter());
    g_scene = scene;
    return 0;
}

// Renders rows [y0,y1) of a width x height frame with `samples` spp on `threads` OpenMP threads.
// rgb_out (may be NULL): (y1-y0)*width*3 bytes.  Returns 0; *rays_out = trace_ray calls, *seconds_out = wall time.
int cpuref_render(int width, int height, int samples, int y0, int y1, int threads,
                  unsigned long long* rays_out, double* seconds_out, unsigned char* rgb_out)
{
    if (!g_scene || width < 1 || height < 1 || samples < 1 || y0 < 0 || y1 > height || y0 >= y1) return -1;
    if (threads < 1) threads = 1;
    g_counts.assign((size_t)threads, Counter{0});
    int saved = dup(1); int devnull = open("/dev/null", O_WRONLY);   // Camera ctor prints its basis (camera.cpp:29-33)
    fflush(stdout); dup2(devnull, 1);
    Camera camera(glm::dvec3(-2, -5, 2.5), glm::dvec3(0, 0, 0), width, height);    // main.cpp:22
    fflush(stdout); dup2(saved, 1); close(saved); close(devnull);
    const double samples_recp = 1. / samples;
    auto t0 = std::chrono::steady_clock::now();
#pragma omp parallel for schedule(dynamic, 1) num_threads(threads)
    for (int y = y0; y < y1; y++) {
        unsigned short Xi[3] = {0, 0, (unsigned short)(y * y * y)};                // renderer.cpp:205
        for (int x = 0; x < width; x++) {
            glm::dvec3 col = glm::dvec3();
            for (int a = 0; a < samples; a++) {
                Ray ray = camera.get_ray(x, y, a > 0, Xi);                          // :211
                col = col + g_scene->trace_ray(ray, 0, Xi);                         // :212
            }
            if (rgb_out) {
                unsigned char* px = rgb_out + 3 * ((size_t)(y - y0) * width + x);
                px[0] = (unsigned char)toInt((col * samples_recp).x);
                px[1] = (unsigned char)toInt((col * samples_recp).y);
                px[2] = (unsigned char)toInt((col * samples_recp).z);
            }
        }
    }
    auto t1 = std::chrono::steady_clock::now();
    unsigned long long total = 0;
    for (auto& c : g_counts) total += c.n;
    if (rays_out) *rays_out = total;
    if (seconds_out) *seconds_out = std::chrono::duration<double>(t1 - t0).count();
    return 0;
}

} // extern "C"
