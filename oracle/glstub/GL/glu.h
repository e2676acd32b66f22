/* Headless no-op GLU stub (TEST INFRASTRUCTURE, oracle build only); see gl.h. */
#ifndef GI_ORACLE_GLSTUB_GLU_H
#define GI_ORACLE_GLSTUB_GLU_H
#include "gl.h"

struct GLUquadric { int unused; };
typedef GLUquadric GLUquadricObj;

template <class... A> static inline GLUquadric* gluNewQuadric(A...) { static GLUquadric q; return &q; }
GI_GLSTUB_FN(gluDeleteQuadric) GI_GLSTUB_FN(gluQuadricTexture) GI_GLSTUB_FN(gluQuadricNormals)
GI_GLSTUB_FN(gluQuadricDrawStyle) GI_GLSTUB_FN(gluQuadricOrientation) GI_GLSTUB_FN(gluCylinder)
GI_GLSTUB_FN(gluOrtho2D) GI_GLSTUB_FN(gluDisk) GI_GLSTUB_FN(gluSphere) GI_GLSTUB_FN(gluPerspective)
GI_GLSTUB_FN_RET(gluBuild2DMipmaps, int, 0) GI_GLSTUB_FN(gluLookAt)
GI_GLSTUB_FN_RET(gluProject, int, 0) GI_GLSTUB_FN_RET(gluUnProject, int, 0)

enum { GLU_FILL = 0x5000, GLU_FLAT, GLU_INSIDE, GLU_NONE, GLU_OUTSIDE, GLU_SILHOUETTE, GLU_SMOOTH,
       GLU_LINE, GLU_POINT };

#endif
