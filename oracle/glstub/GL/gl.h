/* Headless no-op OpenGL stub (TEST INFRASTRUCTURE, oracle build only).
 *
 * The reference's GAPS headers include <GL/gl.h>/<GL/glu.h> unconditionally
 * (RNBasics/RNGrfx.h:41-42) although `photonmap` never calls GL at run time;
 * only the Draw()/Outline() methods mention these names.  No graphics stack
 * exists on this box, so every entry point is a variadic no-op template and
 * every enum is a distinct integer constant.  Written from scratch: contains no
 * Khronos header text.
 */
#ifndef GI_ORACLE_GLSTUB_GL_H
#define GI_ORACLE_GLSTUB_GL_H

typedef unsigned int GLenum;
typedef unsigned int GLuint;
typedef int GLint;
typedef int GLsizei;
typedef unsigned char GLubyte;
typedef unsigned char GLboolean;
typedef float GLfloat;
typedef double GLdouble;
typedef void GLvoid;
typedef unsigned int GLbitfield;
typedef short GLshort;
typedef unsigned short GLushort;
typedef signed char GLbyte;
typedef float GLclampf;
typedef double GLclampd;

#define GI_GLSTUB_FN(name) template <class... A> static inline void name(A...) {}
#define GI_GLSTUB_FN_RET(name, T, v) template <class... A> static inline T name(A...) { return v; }

GI_GLSTUB_FN(glLightf) GI_GLSTUB_FN(glMatrixMode) GI_GLSTUB_FN(glLightfv) GI_GLSTUB_FN(glEnable)
GI_GLSTUB_FN(glEnd) GI_GLSTUB_FN(glBegin) GI_GLSTUB_FN(glTexParameteri) GI_GLSTUB_FN(glPushMatrix)
GI_GLSTUB_FN(glPopMatrix) GI_GLSTUB_FN(glDisable) GI_GLSTUB_FN(glLoadIdentity) GI_GLSTUB_FN(glBindTexture)
GI_GLSTUB_FN(glDrawPixels) GI_GLSTUB_FN(glReadPixels) GI_GLSTUB_FN(glMaterialfv) GI_GLSTUB_FN(glFrontFace)
GI_GLSTUB_FN(glCallLists) GI_GLSTUB_FN(glBindBuffer) GI_GLSTUB_FN(glTexImage2D) GI_GLSTUB_FN(glTexEnvi)
GI_GLSTUB_FN(glRasterPos2i) GI_GLSTUB_FN(glMultMatrixd) GI_GLSTUB_FN(glGenTextures) GI_GLSTUB_FN(glColor3ub)
GI_GLSTUB_FN(glNewList) GI_GLSTUB_FN(glLoadMatrixd) GI_GLSTUB_FN(glLightModeli) GI_GLSTUB_FN(glGetIntegerv)
GI_GLSTUB_FN_RET(glGenLists, GLuint, 0u) GI_GLSTUB_FN(glGenBuffers) GI_GLSTUB_FN(glEndList)
GI_GLSTUB_FN(glEnableClientState) GI_GLSTUB_FN(glDisableClientState) GI_GLSTUB_FN(glDepthMask)
GI_GLSTUB_FN(glDeleteTextures) GI_GLSTUB_FN(glDeleteLists) GI_GLSTUB_FN(glDeleteBuffers) GI_GLSTUB_FN(glCallList)
GI_GLSTUB_FN(glBufferData) GI_GLSTUB_FN(glBlendFunc) GI_GLSTUB_FN(glViewport) GI_GLSTUB_FN(glVertexPointer)
GI_GLSTUB_FN(glVertex3fv) GI_GLSTUB_FN(glVertex3f) GI_GLSTUB_FN(glVertex3dv) GI_GLSTUB_FN(glVertex3d)
GI_GLSTUB_FN(glVertex2fv) GI_GLSTUB_FN(glVertex2f) GI_GLSTUB_FN(glVertex2dv) GI_GLSTUB_FN(glVertex2d)
GI_GLSTUB_FN(glTranslated) GI_GLSTUB_FN(glTexCoordPointer) GI_GLSTUB_FN(glTexCoord3f) GI_GLSTUB_FN(glTexCoord3d)
GI_GLSTUB_FN(glTexCoord2fv) GI_GLSTUB_FN(glTexCoord2f) GI_GLSTUB_FN(glTexCoord2dv) GI_GLSTUB_FN(glTexCoord2d)
GI_GLSTUB_FN(glRasterPos3f) GI_GLSTUB_FN(glRasterPos3d) GI_GLSTUB_FN(glRasterPos2f) GI_GLSTUB_FN(glRasterPos2d)
GI_GLSTUB_FN(glNormal3fv) GI_GLSTUB_FN(glNormal3f) GI_GLSTUB_FN(glNormal3dv) GI_GLSTUB_FN(glNormal3d)
GI_GLSTUB_FN(glMultMatrixf) GI_GLSTUB_FN(glMaterialf) GI_GLSTUB_FN(glLoadMatrixf) GI_GLSTUB_FN(glGetDoublev)
GI_GLSTUB_FN(glDrawElements) GI_GLSTUB_FN(glColor3fv) GI_GLSTUB_FN(glColor3f) GI_GLSTUB_FN(glColor3dv)
GI_GLSTUB_FN(glColor3d) GI_GLSTUB_FN(glAlphaFunc) GI_GLSTUB_FN(glClear) GI_GLSTUB_FN(glClearColor)
GI_GLSTUB_FN(glFlush) GI_GLSTUB_FN(glFinish) GI_GLSTUB_FN(glLineWidth) GI_GLSTUB_FN(glPointSize)
GI_GLSTUB_FN(glScaled) GI_GLSTUB_FN(glRotated) GI_GLSTUB_FN(glPixelStorei) GI_GLSTUB_FN(glPolygonMode)
GI_GLSTUB_FN(glShadeModel) GI_GLSTUB_FN(glDepthFunc) GI_GLSTUB_FN(glCullFace) GI_GLSTUB_FN(glHint)
GI_GLSTUB_FN(glPolygonOffset) GI_GLSTUB_FN(glColorMaterial) GI_GLSTUB_FN(glLightModelfv) GI_GLSTUB_FN(glOrtho)
GI_GLSTUB_FN(glFrustum) GI_GLSTUB_FN(glGetFloatv) GI_GLSTUB_FN(glNormalPointer) GI_GLSTUB_FN(glColorPointer)
GI_GLSTUB_FN(glDrawArrays) GI_GLSTUB_FN(glColor4f) GI_GLSTUB_FN(glColor4d) GI_GLSTUB_FN(glColor4fv)

enum {
  GL_FALSE = 0, GL_TRUE = 1, GL_ZERO = 0, GL_ONE = 1,
  GL_ALPHA_TEST = 0x100, GL_AMBIENT, GL_ARRAY_BUFFER, GL_BLEND, GL_CCW, GL_CLAMP, GL_COMPILE,
  GL_CONSTANT_ATTENUATION, GL_CULL_FACE, GL_CW, GL_DECAL, GL_DIFFUSE, GL_ELEMENT_ARRAY_BUFFER,
  GL_ELEMENT_BUFFER, GL_EMISSION, GL_FLOAT, GL_FRONT, GL_BACK, GL_FRONT_AND_BACK, GL_GREATER,
  GL_LIGHTING, GL_LIGHT_MODEL_TWO_SIDE, GL_LIGHT_MODEL_AMBIENT, GL_LINEAR, GL_LINEAR_ATTENUATION,
  GL_LINEAR_MIPMAP_LINEAR, GL_LINEAR_MIPMAP_NEAREST, GL_LINE_LOOP, GL_LINE_STRIP, GL_LINES, GL_POINTS,
  GL_LUMINANCE, GL_LUMINANCE_ALPHA, GL_MODELVIEW, GL_MODULATE, GL_NEAREST, GL_NEAREST_MIPMAP_LINEAR,
  GL_NEAREST_MIPMAP_NEAREST, GL_ONE_MINUS_SRC_ALPHA, GL_POLYGON, GL_POSITION, GL_PROJECTION,
  GL_PROJECTION_MATRIX, GL_MODELVIEW_MATRIX, GL_QUADRATIC_ATTENUATION, GL_REPEAT, GL_RGB, GL_RGBA,
  GL_SHININESS, GL_SPECULAR, GL_SPOT_CUTOFF, GL_SPOT_DIRECTION, GL_SPOT_EXPONENT, GL_SRC_ALPHA,
  GL_STATIC_DRAW, GL_TEXTURE_2D, GL_TEXTURE_COORD_ARRAY, GL_TEXTURE_ENV, GL_TEXTURE_ENV_MODE,
  GL_TEXTURE_MAG_FILTER, GL_TEXTURE_MIN_FILTER, GL_TEXTURE_WRAP_S, GL_TEXTURE_WRAP_T, GL_TRIANGLES,
  GL_TRIANGLE_STRIP, GL_TRIANGLE_FAN, GL_QUADS, GL_QUAD_STRIP, GL_UNSIGNED_BYTE, GL_UNSIGNED_INT,
  GL_UNSIGNED_SHORT, GL_VERTEX_ARRAY, GL_NORMAL_ARRAY, GL_COLOR_ARRAY, GL_VIEWPORT, GL_DEPTH_TEST,
  GL_COLOR_BUFFER_BIT, GL_DEPTH_BUFFER_BIT, GL_NORMALIZE, GL_FILL, GL_LINE, GL_POINT, GL_SMOOTH, GL_FLAT,
  GL_LIGHT0 = 0x4000, GL_LIGHT1, GL_LIGHT2, GL_LIGHT3, GL_LIGHT4, GL_LIGHT5, GL_LIGHT6, GL_LIGHT7
};

#endif
