/* kpSift3D -- keypoints and descriptors of one NIfTI volume, on the GPU.
 *
 * Command-line twin of the reference's cli/kpSift3D.c:96-163: same options, same call
 * sequence through the public API (read, detect, keep the 100 strongest, save keypoints,
 * extract, save descriptors), same exit codes. Written against include/sift3d.h; the only
 * difference in behaviour is that the image reader works without nifticlib.
 *
 *   kpSift3D --keys keys.csv --desc desc.csv.gz image.nii.gz
 */
#include <getopt.h>
#include <stdio.h>
#include <stdlib.h>

#include "sift3d.h"

enum { OPT_KEYS = 'k', OPT_DESC = 'd', OPT_HELP = 'h' };

static const char usage[] =
    "Usage: kpSift3D [image.nii] \n"
    "\n"
    "Detects SIFT3D keypoints and extracts their descriptors from an image.\n"
    "\n"
    "Example: \n"
    " kpSift3D --keys keys.csv --desc desc.csv image.nii \n"
    "\n"
    "Output options: \n"
    " --keys [filename] \n"
    "       Specifies the output file name for the keypoints. \n"
    "       Supported file formats: .csv, .csv.gz \n"
    " --desc [filename] \n"
    "       Specifies the output file name for the descriptors. \n"
    "       Supported file formats: .csv, .csv.gz \n"
    "At least one of the output options must be specified. \n";

int main(int argc, char *argv[]) {
    static const struct option opts[] = {{"keys", required_argument, NULL, OPT_KEYS},
                                         {"desc", required_argument, NULL, OPT_DESC},
                                         {"help", no_argument, NULL, OPT_HELP},
                                         {0, 0, 0, 0}};
    const char *keys_path = NULL, *desc_path = NULL, *image_path;
    sift3d_detector *det = NULL;
    sift3d_image *image = NULL;
    sift3d_keypoint_store *kp = NULL;
    sift3d_descriptor_store *desc = NULL;
    int status = 1, c;

    opterr = 1;
    while ((c = getopt_long(argc, argv, "", opts, NULL)) != -1) {
        if (c == OPT_KEYS)
            keys_path = optarg;
        else if (c == OPT_DESC)
            desc_path = optarg;
        else if (c == OPT_HELP) {
            puts(usage);
            return 0;
        } else
            return 1;
    }
    if (keys_path == NULL && desc_path == NULL) {
        fprintf(stderr, "No outputs specified.\n");
        return 1;
    }
    if (argc - optind < 1) {
        fprintf(stderr, "Not enough arguments.\n");
        return 1;
    }
    if (argc - optind > 1) {
        fprintf(stderr, "Too many arguments.\n");
        return 1;
    }
    image_path = argv[optind];

    if ((det = sift3d_make_detector()) == NULL) {
        fprintf(stderr, "Failed to initialize SIFT data.\n");
        goto done;
    }
    kp = sift3d_make_keypoint_store();
    desc = sift3d_make_descriptor_store();
    if ((image = sift3d_read_image(image_path)) == NULL) {
        fprintf(stderr, "Could not read image.\n");
        goto done;
    }
    if (sift3d_detect_keypoints(det, image, kp)) {
        fprintf(stderr, "Failed to detect keypoints.\n");
        goto done;
    }
    /* the reference keeps the 100 "strongest" keypoints (kpSift3D.c:122) */
    sift3d_keypoint_store_sort_by_strength(kp, 100);
    if (keys_path != NULL && sift3d_keypoint_store_save(keys_path, kp)) {
        fprintf(stderr, "Failed to write the keypoints to %s\n", keys_path);
        goto done;
    }
    if (desc_path != NULL) {
        if (sift3d_extract_descriptors(det, kp, desc)) {
            fprintf(stderr, "Failed to extract descriptors.\n");
            goto done;
        }
        if (sift3d_descriptor_store_save(desc_path, desc)) {
            fprintf(stderr, "Failed to write the descriptors to %s\n", desc_path);
            goto done;
        }
    }
    status = 0;

done:
    if (desc != NULL)
        sift3d_free_descriptor_store(desc);
    if (kp != NULL)
        sift3d_free_keypoint_store(kp);
    if (det != NULL)
        sift3d_free_detector(det);
    if (image != NULL)
        sift3d_free_image(image);
    return status;
}
