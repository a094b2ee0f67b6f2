"""Wire format of RPN proposals between the alternate-training stages (reference:
rcnn/rpn/generate.py:60-76 writes it, helper/dataset/pascal_voc.py:190-201 reads it).

The file ``<root>/rpn_data/<imdb.name>_rpn.pkl`` is one pickled python list with, per image in imdb order, a
(post, 4) float32 array ``rois[:, 1:] / scale`` -- boxes in ORIGINAL image coordinates, padding rows included.
"""
import os
import pickle

import numpy as np

# cPickle.HIGHEST_PROTOCOL of the reference's python 2; files written here load there and vice versa
PICKLE_PROTOCOL = 2


def proposals_to_boxes(rois, scale):
    """generate.py:60-62 for one image: drop the batch column, undo the test-time scale (float32 / float32)."""
    rois = np.asarray(rois, dtype=np.float32)
    return rois[:, 1:].copy() / np.float32(scale)


def batch_to_box_list(rois, im_info, rows_per_image):
    """The proposal op's batched output (B*rows, 5) + im_info (B,3) -> list of B (rows,4) float32 arrays."""
    rois = np.asarray(rois, dtype=np.float32).reshape(-1, rows_per_image, 5)
    info = np.asarray(im_info, dtype=np.float32).reshape(-1, 3)
    assert rois.shape[0] == info.shape[0]
    return [proposals_to_boxes(rois[b], info[b, 2]) for b in range(rois.shape[0])]


def rpn_file_path(root_path, imdb_name):
    return os.path.join(root_path, 'rpn_data', imdb_name + '_rpn.pkl')


def write_rpn_file(root_path, imdb_name, imdb_boxes):
    """generate.py:69-76.  Returns the path."""
    folder = os.path.join(root_path, 'rpn_data')
    if not os.path.exists(folder):
        os.mkdir(folder)
    path = rpn_file_path(root_path, imdb_name)
    with open(path, 'wb') as f:
        pickle.dump([np.asarray(b, dtype=np.float32) for b in imdb_boxes], f, PICKLE_PROTOCOL)
    print('wrote rpn proposals to {}'.format(path))
    return path


def read_rpn_file(root_path, imdb_name):
    """pascal_voc.py:196-200: the ``box_list`` handed to ``create_roidb_from_box_list``."""
    path = rpn_file_path(root_path, imdb_name)
    assert os.path.exists(path), 'rpn data not found at {}'.format(path)
    with open(path, 'rb') as f:
        return pickle.load(f, encoding='latin1')
