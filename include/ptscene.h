/* ptscene.h — C ABI of libptscene.so: the HOST side of the path tracer (scene construction).
 *
 * This library keeps the reference's scene API and data layout (north star: "the host side stays C++
 * and keeps the reference's scene API and data layout"), re-implemented from scratch so that every
 * output array is BIT-IDENTICAL to what the reference's main.cpp would hand to OpenGL.  All
 * citations are to /root/reference/ImportanceSampling_LowDiscrepancySequence/main.cpp ("M:") —
 * OpenglRayTracing/main.cpp and DisneyBRDF/main.cpp carry the same functions.
 *
 * No CUDA here; the arrays this library produces are what pt_upload_scene (include/pt.h) consumes.
 * Every function returns 0 / a count on success and a negative PTS_E* code on failure; none calls
 * exit() (the reference does, M:176, M:270).
 */
#ifndef PTSCENE_H
#define PTSCENE_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PTS_OK            0
#define PTS_EINVAL       -1
#define PTS_EIO          -2
#define PTS_EFORMAT      -3
#define PTS_ENOMEM       -4
#define PTS_ERANGE       -5   /* an index no longer fits the float encoding (>= 2^24), M:817-818 */

#define PTS_TRI_FLOATS   36   /* Triangle / Triangle_encoded: 12 x vec3, M:44-60 (same memory layout) */
#define PTS_NODE_FLOATS  12   /* BVHNode_encoded: 4 x vec3, M:68-72 */

/* struct Material, M:26-41 — 18 packed floats, the tail of every encoded triangle (texels 6..11) */
typedef struct PtsMaterial {
    float emissive[3];
    float baseColor[3];
    float subsurface, metallic, specular;
    float specularTint, roughness, anisotropic;
    float sheen, sheenTint, clearcoat;
    float clearcoatGloss, IOR, transmission;
} PtsMaterial;

/* the member initialisers of struct Material (baseColor 1, IOR 1, everything else 0) */
void pts_material_default(PtsMaterial* m);

/* getTransformMatrix, M:243-259: translate * rotateX*rotateY*rotateZ (degrees) * scale, column-major */
void pts_get_transform_matrix(const float rotate_deg[3], const float translate[3], const float scale[3],
                              float out16[16]);

/* camera lines of display(), M:671-674: eye on the orbit sphere, cameraRotate = inverse(lookAt(eye,0,+Y)) */
void pts_camera(float upAngle_deg, float rotateAngle_deg, float r, float eye3[3], float cameraRotate16[16]);

/* ---- scene = std::vector<Triangle> + std::vector<BVHNode> --------------------------------------- */
typedef struct pts_scene pts_scene;
pts_scene* pts_scene_create(void);
void pts_scene_destroy(pts_scene* s);

/* readObj, M:262-373: `v` / `f` parser (v, v/vt, v/vt/vn), unit-normalisation by the reference's bbox
   (including its maxy/miny/maxz/minz quirk, M:298-299), model matrix, face or smooth normals.
   Appends to the scene; returns the number of triangles appended. */
int pts_read_obj(pts_scene* s, const char* path, const PtsMaterial* m, const float trans16[16], int smoothNormal);

/* Same pipeline as readObj from in-memory arrays (what readObj does after parsing): used by the
   synthetic mesh generators so that generated meshes are "scaled as readObj would". */
int pts_add_mesh(pts_scene* s, const float* vertices_xyz, int n_vertices, const unsigned* indices, int n_indices,
                 const PtsMaterial* m, const float trans16[16], int smoothNormal);

/* Appends already-built triangles (36 floats each) verbatim. */
int pts_add_triangles(pts_scene* s, const float* tris, int n);

/* Synthetic meshes (seeded, deterministic, no libm):
   displaced geodesic icosphere, frequency n -> 20*n*n triangles, radius 1 + amplitude*fbm(seed);
   unit quad in the xz plane (the missing models/quad.obj: 2 triangles, y = 0, [-0.5,0.5]^2). */
int pts_add_icosphere(pts_scene* s, int frequency, float amplitude, unsigned seed,
                      const PtsMaterial* m, const float trans16[16], int smoothNormal);
int pts_add_quad(pts_scene* s, const PtsMaterial* m, const float trans16[16]);
int pts_write_obj_icosphere(const char* path, int frequency, float amplitude, unsigned seed);

int pts_triangle_count(const pts_scene* s);
int pts_node_count(const pts_scene* s);

/* buildBVH (M:377-428, use_sah = 0) or buildBVHwithSAH (M:431-552, use_sah = 1) over all triangles
   with leaf size n (the reference passes 8), preceded by the dummy node 0 of M:780-786.  Reorders the
   triangles exactly as the reference's std::sort calls would.  n_threads <= 1: sequential;
   > 1: independent subtrees are built as parallel tasks (same output).  Returns nNodes. */
int pts_build_bvh(pts_scene* s, int leaf_n, int use_sah, int n_threads);

/* The encoded arrays of M:793-821, owned by the scene, valid until the next mutating call. */
const float* pts_triangles_encoded(const pts_scene* s);   /* pts_triangle_count() x 36 floats */
const float* pts_nodes_encoded(const pts_scene* s);       /* pts_node_count()     x 12 floats */

/* Binary cache of a built scene (triangles + nodes); host SAH build is minutes at >= 10M triangles. */
int pts_scene_save(const pts_scene* s, const char* path);
int pts_scene_load(pts_scene* s, const char* path);

/* ---- HDR environment ---------------------------------------------------------------------------- */
/* HDRLoader::load, hdrloader.cpp:29-97 (Radiance RGBE, new and old RLE), rows in file order, with the
   `%ld`-into-int header bug fixed.  *data is malloc'ed: release with pts_free. */
int pts_hdr_load(const char* path, int* width, int* height, float** data);
int pts_hdr_write(const char* path, const float* data, int width, int height);   /* flat RGBE writer for tests */
/* calculateHdrCache, M:555-652: out[h*w*3] = (x/w, y/h, pdf) inverse-CDF lookup table, fp32 running sums */
int pts_hdr_cache(const float* hdr, int width, int height, float* out);
/* procedural sky (sun + gradient + ground), deterministic, for runs without an .hdr file */
int pts_hdr_synth(int width, int height, unsigned seed, float* out);
void pts_free(void* p);

/* pass3.fsh:14-24 tone map c/(1+lum/1.5) and a PPM/PFM writer for headless output */
int pts_write_pfm(const char* path, const float* rgba, int width, int height);
int pts_write_ppm_tonemapped(const char* path, const float* rgba, int width, int height);
/* 8-bit PNG (stored deflate blocks, what svpng of Basic/main.cpp:169-190 produces): rows top first, 3 or 4 channels.
   Feed it pt_readback_display's bytes to get the image the reference's window shows. */
int pts_write_png(const char* path, const unsigned char* pixels, int width, int height, int channels);

const char* pts_last_error(void);

#ifdef __cplusplus
}
#endif
#endif /* PTSCENE_H */
