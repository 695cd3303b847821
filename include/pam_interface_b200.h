/* pam_interface_b200.h -- the GCM-facing host registry of PAM (pam_core/pam_interface): exactly the extern "C" symbols of
 * pam_core/pam_interface/pam_interface_extern_c.cpp:5-274 that pam_fortran_interface.F90 binds, exported by
 * libawfl_b200.so so that an E3SM-MMF build links this library in place of the reference's pam_interface objects
 * (SURVEY.md 8f row 3). Host-side data only, one registry per host thread (pam_interface.h:9-32). <T> = bool, int,
 * float, double. Arguments written `T &x` in the reference are C++ references (= pointers in the ABI, which is how the
 * Fortran interface passes them); this header spells them as C pointers.
 * Errors: message on stderr, then a C++ exception of type std::string (pam_core/pam_const.h:249-252). */
#ifndef PAM_INTERFACE_B200_H
#define PAM_INTERFACE_B200_H

#ifdef __cplusplus
#define PAMI_REF(T) T &
extern "C" {
#else
#include <stdbool.h>
#define PAMI_REF(T) T *
#endif

void pam_interface_finalize(void);                                                   /* :5 */

#define PAMI_DECLARE(SUFFIX, T)                                                                                          \
  void pam_interface_mirror_array_readonly_##SUFFIX(char const *name, char const *desc, int *dims, int ndims, T *ptr);   \
  void pam_interface_mirror_array_readwrite_##SUFFIX(char const *name, char const *desc, int *dims, int ndims, T *ptr);  \
  void pam_interface_create_array_##SUFFIX(char const *name, char const *desc, int *dims, int ndims);                    \
  void pam_interface_get_array_##SUFFIX(char const *name, PAMI_REF(T *) ptr, int *dims, int ndims);                      \
  void pam_interface_set_option_##SUFFIX(char const *name, T value);                                                     \
  void pam_interface_get_option_##SUFFIX(char const *name, PAMI_REF(T) value);
PAMI_DECLARE(bool, bool)        /* :10, :60, :165, :221, :111, :128 */
PAMI_DECLARE(int, int)          /* :22, :72, :177, :231, :114, :131 */
PAMI_DECLARE(float, float)      /* :34, :84, :189, :241, :120, :142 */
PAMI_DECLARE(double, double)    /* :46, :96, :201, :251, :123, :145 */

void pam_interface_set_option_string(char const *name, char const *value);           /* :117 */
void pam_interface_get_option_stringlen(char const *name, PAMI_REF(int) len);        /* :134 */
void pam_interface_get_option_string(char const *name, char *value);                 /* :138 */
void pam_interface_option_exists(char const *name, PAMI_REF(bool) exists);           /* :150 */
void pam_interface_remove_option(char const *name);                                  /* :155 */
void pam_interface_make_readonly(char const *name);                                  /* :160 */
void pam_interface_destroy_array(char const *name);                                  /* :215 */
void pam_interface_array_exists(char const *name, PAMI_REF(bool) exists);            /* :263 */
void pam_interface_register_dimension(char const *name, int len);                    /* :268 */

/* B200 additions: carry a registered host array to / from device memory (byte count = product of dims x element size). */
int pam_interface_b200_array_bytes(char const *name, unsigned long *bytes);
int pam_interface_b200_copy_to_device(char const *name, void *device_ptr);
int pam_interface_b200_copy_from_device(char const *name, const void *device_ptr);

#ifdef __cplusplus
}
#endif
#endif
