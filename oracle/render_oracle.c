/* TEST INFRASTRUCTURE -- CPU restatement of the headless consumer of the vertex stream
 * (SURVEY.md 8f rank 3): what the reference does with a ChunkMesh between `mesh_for_chunk` and the
 * rasteriser.  Never linked into the product library.
 *
 *  - chunk culling: mimas-client/client.rs:679-696 (viewing range on the chunk ORIGIN, then a
 *    sphere-vs-frustum test on the chunk centre with r = 16 * sqrt(3))
 *  - the frustum: `Frustum::from_modelview_and_projection_2d` / `sphere_intersecting` of the
 *    third-party crate frustum_query 0.1.2 (Cargo.lock:690-693; sources NOT under /root/reference).
 *    Restated from its published algorithm, the classic clip-matrix plane extraction
 *    (clip = modelview * projection on column-major arrays, six planes, each normalised by the length
 *    of its normal; a sphere is outside iff a*x + b*y + c*z + d <= -r for some plane).
 *  - the vertex stage: mimas-client/vertex-shader.glsl:18-27.
 *
 * PARITY UNPINNED: the reference has no tests or golden vectors for any of this, the GL path is
 * unavailable on this box and frustum_query is absent, so the order of the float operations below is
 * this file's choice (every product and sum separately rounded, left to right, no FMA).  The CUDA
 * kernels (mimas_b200/csrc/mm_render.cu) are held bit-exact to THIS file.
 *
 * Matrices are `[[f32; 4]; 4]` as the reference hands them to glium (client.rs:599-600, 1388-1400):
 * column-major, m[4*col + row]. */
#include <math.h>
#include <stdint.h>
#include <string.h>

typedef struct ro_camera {
	float vmatrix[16];
	float pmatrix[16];
	float pos[3];
	float viewing_range;
} ro_camera;

typedef struct ro_vertex {  /* = mm_vertex */
	uint16_t tex_ind, pad0;
	float tex_pos[2], position[3], normal[3];
} ro_vertex;

typedef struct ro_shaded {  /* = mm_shaded_vertex: the vertex shader's outputs */
	float clip[4];       /* gl_Position */
	float vposition[4];
	float vlamb;
	float vtex_pos[2];
	uint32_t vtex_ind;
} ro_shaded;

/* frustum_query: planes right, left, bottom, top, far, near */
void ro_frustum(const float *modl, const float *proj, float planes[6][4]) {
	float clip[16];
	for (int r = 0; r < 4; r++)
		for (int c = 0; c < 4; c++)
			clip[4 * r + c] = modl[4 * r + 0] * proj[0 + c] + modl[4 * r + 1] * proj[4 + c] +
			                  modl[4 * r + 2] * proj[8 + c] + modl[4 * r + 3] * proj[12 + c];
	static const int comp[6] = {0, 0, 1, 1, 2, 2};
	static const float sign[6] = {-1.0f, 1.0f, 1.0f, -1.0f, -1.0f, 1.0f};
	for (int p = 0; p < 6; p++) {
		float v[4];
		for (int k = 0; k < 4; k++) v[k] = clip[4 * k + 3] + sign[p] * clip[4 * k + comp[p]];
		const float t = sqrtf(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
		for (int k = 0; k < 4; k++) planes[p][k] = v[k] / t;
	}
}

int ro_sphere_intersecting(const float planes[6][4], float x, float y, float z, float r) {
	for (int p = 0; p < 6; p++)
		if (planes[p][0] * x + planes[p][1] * y + planes[p][2] * z + planes[p][3] <= -r) return 0;
	return 1;
}

/* client.rs:679-696 for chunk origins offs[n][3] (blocks); writes the indices of the chunks to draw in
 * chunk order and returns how many. */
uint32_t ro_cull_chunks(const int64_t *offs, uint32_t n, const ro_camera *cam, uint32_t *draw) {
	float planes[6][4];
	ro_frustum(cam->vmatrix, cam->pmatrix, planes);
	const float r = 16.0f * sqrtf(3.0f);
	uint32_t k = 0;
	for (uint32_t i = 0; i < n; i++) {
		const float dx = (float)offs[3 * i] - cam->pos[0], dy = (float)offs[3 * i + 1] - cam->pos[1],
		            dz = (float)offs[3 * i + 2] - cam->pos[2];
		if (sqrtf(dx * dx + dy * dy + dz * dz) > cam->viewing_range) continue;
		const float cx = (float)(offs[3 * i] + 8), cy = (float)(offs[3 * i + 1] + 8), cz = (float)(offs[3 * i + 2] + 8);
		if (!ro_sphere_intersecting(planes, cx, cy, cz, r)) continue;
		draw[k++] = i;
	}
	return k;
}

static void mat_vec(const float *m, const float *v, float *out) {
	for (int r = 0; r < 4; r++) out[r] = m[r] * v[0] + m[4 + r] * v[1] + m[8 + r] * v[2] + m[12 + r] * v[3];
}

/* vertex-shader.glsl:18-27 */
void ro_vertex_stage(const ro_vertex *in, uint64_t n, const ro_camera *cam, ro_shaded *out) {
	const float l = 1.0f / sqrtf(2.0f);  /* normalize(vec3(0, -1, -1)) = (0, -l, -l); dir_light_b = (0, l, l) */
	for (uint64_t i = 0; i < n; i++) {
		const float p[4] = {in[i].position[0], in[i].position[1], in[i].position[2], 1.0f};
		ro_shaded o;
		memset(&o, 0, sizeof(o));
		mat_vec(cam->vmatrix, p, o.vposition);
		mat_vec(cam->pmatrix, o.vposition, o.clip);
		const float *nn = in[i].normal;
		const float len = sqrtf(nn[0] * nn[0] + nn[1] * nn[1] + nn[2] * nn[2]);
		const float nx = nn[0] / len, ny = nn[1] / len, nz = nn[2] / len;
		const float da = 0.0f * nx + (-l) * ny + (-l) * nz, db = 0.0f * nx + l * ny + l * nz;
		o.vlamb = fmaxf(da, 0.2f) + fmaxf(db, 0.2f);
		o.vtex_pos[0] = in[i].tex_pos[0];
		o.vtex_pos[1] = in[i].tex_pos[1];
		o.vtex_ind = in[i].tex_ind;
		out[i] = o;
	}
}
