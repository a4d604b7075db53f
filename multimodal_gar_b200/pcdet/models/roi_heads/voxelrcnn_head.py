"""Host-side mirror of VoxelRCNNHead as MGAR-net ships it (/root/reference/pcdet/models/roi_heads/voxelrcnn_head.py:8-239;
SURVEY 8 f1): per-person RoI grid pooling over the multi-scale voxel features + the shared FC stack.  This is the
LiDAR path `Multimodal_cfg/mil3.yaml:105-135` executes: 6x6x6 grid points per person box, three feature scales,
voxel query + grouping + shared MLP + max (NeighborVoxelSAModuleMSG on libmgar_lidar.so).

Same constructor arguments, sub-module names (`roi_grid_pool_layers.{k}.*`, `shared_fc_layer.*` - checkpoints load
strictly), batch_dict keys and method names.  Differences: the class derives from nn.Module (the reference's
RoIHeadTemplate only contributes loss / target-assignment code that the shipped forward never calls), the sparse
tensors may be anything with `.indices` (V,4) [batch,z,y,x], `.features` (V,C), `.spatial_shape`, `.batch_size`
(spconv is not required), and the two per-frame Python loops (:108-110, :135-137) are batched tensor ops.
"""
import torch
import torch.nn as nn

from ...ops.pointnet2.pointnet2_stack import voxel_pool_modules as voxelpool_stack_modules
from ...utils import common_utils


class SparseFeatures:
    """The four attributes of a spconv SparseConvTensor that roi_grid_pool reads."""

    def __init__(self, features, indices, spatial_shape, batch_size):
        self.features, self.indices, self.spatial_shape, self.batch_size = features, indices, list(spatial_shape), batch_size


class VoxelRCNNHead(nn.Module):
    def __init__(self, backbone_channels, model_cfg, point_cloud_range, voxel_size, num_class=1, **kwargs):
        super().__init__()
        self.num_class = num_class
        self.model_cfg = model_cfg
        self.pool_cfg = model_cfg.ROI_GRID_POOL
        layer_cfg = self.pool_cfg.POOL_LAYERS
        self.point_cloud_range = point_cloud_range
        self.voxel_size = voxel_size

        c_out = 0
        self.roi_grid_pool_layers = nn.ModuleList()
        for src_name in self.pool_cfg.FEATURES_SOURCE:
            mlps = layer_cfg[src_name].MLPS
            for k in range(len(mlps)):                                   # in place, as the reference does (:22-23)
                mlps[k] = [backbone_channels[src_name]] + mlps[k]
            self.roi_grid_pool_layers.append(voxelpool_stack_modules.NeighborVoxelSAModuleMSG(
                query_ranges=layer_cfg[src_name].QUERY_RANGES, nsamples=layer_cfg[src_name].NSAMPLE,
                radii=layer_cfg[src_name].POOL_RADIUS, mlps=mlps, pool_method=layer_cfg[src_name].POOL_METHOD))
            c_out += sum(x[-1] for x in mlps)

        grid = self.pool_cfg.GRID_SIZE
        pre_channel = grid * grid * grid * c_out
        shared_fc_list = []
        for k in range(len(model_cfg.SHARED_FC)):
            shared_fc_list.extend([nn.Linear(pre_channel, model_cfg.SHARED_FC[k], bias=False),
                                   nn.BatchNorm1d(model_cfg.SHARED_FC[k]), nn.ReLU(inplace=True)])
            pre_channel = model_cfg.SHARED_FC[k]
            if k != len(model_cfg.SHARED_FC) - 1 and model_cfg.DP_RATIO > 0:
                shared_fc_list.append(nn.Dropout(model_cfg.DP_RATIO))
        self.shared_fc_layer = nn.Sequential(*shared_fc_list)
        self.init_weights()

    def init_weights(self):
        """reference :50-62: xavier_normal on the shared FC linears (no biases there)."""
        for m in self.shared_fc_layer.modules():
            if isinstance(m, nn.Linear):
                nn.init.xavier_normal_(m.weight)
                if m.bias is not None:
                    nn.init.constant_(m.bias, 0)

    def roi_grid_pool(self, batch_dict):
        """reference :78-164.  batch_dict: gt_boxes (B, num_rois, 7+), batch_size, multi_scale_3d_strides {src: int},
        multi_scale_3d_features {src: sparse tensor} (or ..._post with with_voxel_feature_transform)
        -> (B*num_rois, GRID_SIZE^3, sum C_out)."""
        rois = batch_dict['gt_boxes']
        batch_size = batch_dict['batch_size']
        key = 'multi_scale_3d_features_post' if batch_dict.get('with_voxel_feature_transform', False) \
            else 'multi_scale_3d_features'
        grid = self.pool_cfg.GRID_SIZE

        roi_grid_xyz, _ = self.get_global_grid_points_of_roi(rois, grid_size=grid)
        roi_grid_xyz = roi_grid_xyz.view(batch_size, -1, 3)                # (B, num_rois * grid^3, 3)
        # voxel coordinates of the grid points at stride 1, as floats: float floor division like the reference (:101-105)
        roi_grid_coords = torch.cat([(roi_grid_xyz[:, :, a:a + 1] - self.point_cloud_range[a]) // self.voxel_size[a]
                                     for a in range(3)], dim=-1)
        per_frame = roi_grid_coords.shape[1]
        batch_idx = torch.arange(batch_size, device=rois.device, dtype=rois.dtype).view(-1, 1, 1).expand(-1, per_frame, 1)
        roi_grid_batch_cnt = torch.full((batch_size,), per_frame, dtype=torch.int32, device=rois.device)
        new_xyz = roi_grid_xyz.contiguous().view(-1, 3)

        pooled_features_list = []
        for k, src_name in enumerate(self.pool_cfg.FEATURES_SOURCE):
            cur_stride = batch_dict['multi_scale_3d_strides'][src_name]
            cur_sp_tensors = batch_dict[key][src_name]
            cur_coords = cur_sp_tensors.indices
            cur_voxel_xyz = common_utils.get_voxel_centers(cur_coords[:, 1:4], downsample_times=cur_stride,
                                                           voxel_size=self.voxel_size,
                                                           point_cloud_range=self.point_cloud_range)
            cur_voxel_xyz_batch_cnt = torch.bincount(cur_coords[:, 0].long(), minlength=batch_size)[:batch_size].int()
            v2p_ind_tensor = common_utils.generate_voxel2pinds(cur_sp_tensors)
            cur_roi_grid_coords = torch.cat([batch_idx, roi_grid_coords // cur_stride], dim=-1).int()   # [b, x, y, z]
            pooled_features = self.roi_grid_pool_layers[k](
                xyz=cur_voxel_xyz.contiguous(), xyz_batch_cnt=cur_voxel_xyz_batch_cnt, new_xyz=new_xyz,
                new_xyz_batch_cnt=roi_grid_batch_cnt, new_coords=cur_roi_grid_coords.contiguous().view(-1, 4),
                features=cur_sp_tensors.features.contiguous(), voxel2point_indices=v2p_ind_tensor)
            pooled_features_list.append(pooled_features.view(-1, grid ** 3, pooled_features.shape[-1]))
        return torch.cat(pooled_features_list, dim=-1)

    def get_global_grid_points_of_roi(self, rois, grid_size):
        """reference :167-178: grid points of every box in the LiDAR frame, and in the box frame."""
        rois = rois.view(-1, rois.shape[-1])
        local_roi_grid_points = self.get_dense_grid_points(rois, rois.shape[0], grid_size)
        global_roi_grid_points = common_utils.rotate_points_along_z(local_roi_grid_points.clone(), rois[:, 6]).squeeze(dim=1)
        global_roi_grid_points += rois[:, 0:3].clone().unsqueeze(dim=1)
        return global_roi_grid_points, local_roi_grid_points

    @staticmethod
    def get_dense_grid_points(rois, batch_size_rcnn, grid_size):
        """reference :180-189: cell centres of a grid_size^3 lattice spanning each box, (x,y,z) index order."""
        dense_idx = rois.new_ones((grid_size, grid_size, grid_size)).nonzero()
        dense_idx = dense_idx.repeat(batch_size_rcnn, 1, 1).float()
        local_roi_size = rois.view(batch_size_rcnn, -1)[:, 3:6]
        return (dense_idx + 0.5) / grid_size * local_roi_size.unsqueeze(dim=1) - (local_roi_size.unsqueeze(dim=1) / 2)

    def forward(self, batch_dict):
        """reference :191-239 as shipped: pooled_features (B*num_rois, grid^3, C) and shared_feature (B*num_rois, FC)."""
        pooled_features = self.roi_grid_pool(batch_dict)
        batch_dict['pooled_features'] = pooled_features
        batch_dict['shared_feature'] = self._shared_fc(pooled_features.view(pooled_features.size(0), -1))
        return batch_dict

    def _shared_fc(self, x):
        """[Linear (no bias), BatchNorm1d, ReLU(, Dropout)] x len(SHARED_FC).  Eval mode without autograd: BatchNorm folded
        into the weight rows and every layer one tcgen05 GEMM with bias + ReLU epilogue (csrc/gemm_tc.cu, 3xTF32);
        otherwise the reference's nn.Sequential."""
        from .... import tc_gemm
        layers = [m for m in self.shared_fc_layer if isinstance(m, (nn.Linear, nn.BatchNorm1d))]
        pairs = list(zip(layers[0::2], layers[1::2]))
        if self.training or torch.is_grad_enabled() or not all(
                tc_gemm.usable(x, lin.out_features, lin.in_features) for lin, _ in pairs):
            return self.shared_fc_layer(x)
        tensors = [t for lin, bn in pairs for t in (lin.weight, bn.weight, bn.bias, bn.running_mean, bn.running_var)]
        key = tuple((t.data_ptr(), t._version, str(t.device)) for t in tensors)
        if getattr(self, '_fc_key', None) != key:
            self._fc_folded = []
            for lin, bn in pairs:
                scale = bn.weight.detach() / torch.sqrt(bn.running_var.detach() + bn.eps)
                self._fc_folded.append(((lin.weight.detach() * scale.unsqueeze(1)).contiguous(),
                                        (bn.bias.detach() - bn.running_mean.detach() * scale).contiguous()))
            self._fc_key = key
        for w, b in self._fc_folded:
            x = tc_gemm.linear(x, w, b, relu=True)
        return x
